#!/usr/bin/env python
"""Generates the committed fixtures under tests/golden/ (run in the authoring container; the GPU box only reads them).

None of these are produced by anything under oracle/ or llm_fp8_b200/ — they come from code the oracle must agree with:

  fp8_cast_all_bf16.npz   torch's own float8 casts (the library the north star names as the byte standard:
                          `x.float().mul(scale).clamp(-fmax, fmax).to(torch.float8_e4m3fn | e5m2)`) of EVERY bf16 bit
                          pattern at three scales.  Pins oracle/oracle_c.c (integer bit-level) and the CUDA casts.
  thesis_encodings.json   the three valid worked encodings of the reference (thesis/chapters/c2/c2_quantization.tex:100,102,103).
  mx_blocks.npz           1x32 blocks across the fp32 range (scale-step boundaries, zero, denormal, max finite) with the E8M0
                          byte and the 32 e4m3 bytes the north star's MXFP8 rule gives, computed with numpy float32 arithmetic
                          (division, frexp, ldexp) and torch's e4m3 cast.  Pins both oracle restatements and the CUDA quantiser.
  hf_bf16_loss.json       loss trajectory of the reference's CPU-runnable path — the plain HF bf16 LlamaForCausalLM it
                          builds without --use_te (train_fp8.py:118-124) — on a seeded tiny Llama config and seeded
                          synthetic tokens, AdamW on bf16 params (train_fp8.py:200-204; te_llama.py:107).  The loss-band
                          GPU test compares the FP8 decoder swap against this trajectory.
"""
import json
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
SCALES = [1.0, 13.0, 1.0 / 3.0]
TINY = dict(hidden_size=256, intermediate_size=512, num_attention_heads=4, num_key_value_heads=2, num_hidden_layers=2,
            vocab_size=512, max_position_embeddings=128, rms_norm_eps=1e-5, tie_word_embeddings=True)
LOSS_CFG = dict(seed=1234, batch=4, seq=64, steps=8, lr=1e-3)


def all_bf16():
    return torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)


def torch_sat_cast(x_bf16, scale, dtype):
    fmax = torch.finfo(dtype).max
    v = x_bf16.float() * torch.tensor(scale, dtype=torch.float32)
    return v.clamp(-fmax, fmax).to(dtype).view(torch.uint8).numpy()


def make_casts():
    x = all_bf16()
    out = {"scales": np.array(SCALES, dtype=np.float32)}
    for i, s in enumerate(SCALES):
        out[f"e4m3_{i}"] = torch_sat_cast(x, np.float32(s).item(), torch.float8_e4m3fn)
        out[f"e5m2_{i}"] = torch_sat_cast(x, np.float32(s).item(), torch.float8_e5m2)
    np.savez_compressed(os.path.join(HERE, "fp8_cast_all_bf16.npz"), **out)


def make_thesis():
    enc = [
        {"value": 0.2979, "format": "e4m3", "decoded": 0.3125, "exp_field": 5, "mantissa": 2, "src": "thesis/chapters/c2/c2_quantization.tex:100"},
        {"value": 0.2979, "format": "e5m2", "decoded": 0.3125, "exp_field": 13, "mantissa": 1, "src": "thesis/chapters/c2/c2_quantization.tex:102"},
        {"value": 500.0, "format": "e5m2", "decoded": 512.0, "exp_field": 24, "mantissa": 0, "src": "thesis/chapters/c2/c2_quantization.tex:103"},
    ]
    json.dump(enc, open(os.path.join(HERE, "thesis_encodings.json"), "w"), indent=1)


def hf_bf16_losses(tiny=TINY, cfg=LOSS_CFG):
    """The reference's non-TE model path on CPU; also imported by tests/test_golden.py to re-check the fixture."""
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(cfg["seed"])
    model = LlamaForCausalLM(LlamaConfig(**tiny)).to(torch.bfloat16)
    g = torch.Generator().manual_seed(cfg["seed"] + 1)
    tokens = torch.randint(0, tiny["vocab_size"], (cfg["batch"], cfg["seq"]), generator=g)
    opt = torch.optim.AdamW(model.parameters(), lr=cfg["lr"])
    losses = []
    for _ in range(cfg["steps"]):
        out = model(input_ids=tokens, labels=tokens)
        out.loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        losses.append(float(out.loss))
    return losses


def make_loss():
    json.dump({"tiny_config": TINY, "run": LOSS_CFG, "losses": hf_bf16_losses(), "torch": torch.__version__,
               "how": "LlamaForCausalLM(LlamaConfig(**tiny_config)).bfloat16() on CPU, AdamW(lr), same batch every step, labels = input_ids"},
              open(os.path.join(HERE, "hf_bf16_loss.json"), "w"), indent=1)


def mx_block_inputs():
    """[n_blocks, 32] fp32: magnitudes across the whole fp32 range, block maxima on and one ulp either side of the 448 * 2^k
    boundaries where the E8M0 scale steps, zero / denormal / max-finite blocks, and bf16-representable blocks (what the plain
    quantiser sees; the RMSNorm-fused one sees fp32)."""
    g = torch.Generator().manual_seed(77)
    blocks = []
    for k in range(-120, 121, 3):
        blocks.append(torch.randn(32, generator=g) * 2.0 ** k)
    for k in (-100, -20, -9, -8, -1, 0, 1, 7, 8, 40, 100):
        base = np.float32(448.0) * np.float32(2.0) ** np.float32(k)
        for amax in (np.nextafter(base, np.float32(0)), base, np.nextafter(base, np.float32(np.inf))):
            b = torch.randn(32, generator=g) * float(amax) * 0.25
            b = b.clamp(-float(amax), float(amax))
            b[int(torch.randint(0, 32, (1,), generator=g))] = float(amax) * (1 if k % 2 else -1)
            blocks.append(b)
    blocks.append(torch.zeros(32))
    z = torch.zeros(32)
    z[5] = 1e-42                                             # denormal maximum
    blocks.append(z)
    z = torch.randn(32, generator=g) * 1e30
    z[0] = 3.3895313892515355e38                             # largest bf16 value
    blocks.append(z)
    for _ in range(16):
        blocks.append((torch.randn(32, generator=g) * torch.exp(2 * torch.randn(32, generator=g))).bfloat16().float())
    x = torch.stack(blocks).float()
    return torch.cat((x, x.bfloat16().float()))              # every block also rounded to bf16 (the plain quantiser's input type)


def mx_block_expected(x):
    """The north star's MXFP8 rule restated with numpy float32 arithmetic and torch's e4m3 cast — no code shared with oracle/:
    v = amax / 448 in fp32; E8M0 = the smallest power of two >= v (exponent + 127, clamped to [0, 254]; 0 for v = 0);
    data = satfinite RNE e4m3 of x * 2^(127 - E8M0), the product rounded to fp32."""
    xn = x.numpy().astype(np.float32)
    amax = np.abs(xn).max(axis=1)
    v = amax / np.float32(448.0)
    m, ex = np.frexp(v)                                       # v = m * 2^ex, m in [0.5, 1)
    k = np.where(m == np.float32(0.5), ex - 1, ex)
    e = np.clip(k + 127, 0, 254)
    e = np.where(v == 0, 0, e).astype(np.int32)
    inv = np.ldexp(np.float32(1.0), 127 - e).astype(np.float32)
    y = (xn * inv[:, None]).astype(np.float32)
    q = torch.from_numpy(y).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    return e.astype(np.uint8), q


def make_mx_blocks():
    x = mx_block_inputs()
    e, q = mx_block_expected(x)
    np.savez_compressed(os.path.join(HERE, "mx_blocks.npz"), x=x.numpy(), e8m0=e, q=q)


if __name__ == "__main__":
    make_casts()
    make_thesis()
    make_loss()
    make_mx_blocks()
    print("golden fixtures written to", HERE)
