#!/usr/bin/env python
"""Generates the committed fixtures under tests/golden/ (run in the authoring container; the GPU box only reads them).

None of these are produced by anything under oracle/ or llm_fp8_b200/ — they come from code the oracle must agree with:

  fp8_cast_all_bf16.npz   torch's own float8 casts (the library the north star names as the byte standard:
                          `x.float().mul(scale).clamp(-fmax, fmax).to(torch.float8_e4m3fn | e5m2)`) of EVERY bf16 bit
                          pattern at three scales.  Pins oracle/oracle_c.c (integer bit-level) and the CUDA casts.
  thesis_encodings.json   the three valid worked encodings of the reference (thesis/chapters/c2/c2_quantization.tex:100,102,103).
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


if __name__ == "__main__":
    make_casts()
    make_thesis()
    make_loss()
    print("golden fixtures written to", HERE)
