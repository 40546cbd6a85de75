"""Host logic of llm_fp8_b200.lm_head on CPU (the kernels are replaced by torch stand-ins): LMHeadLoss's label shift must be HF's
(logits[..., :-1, :] against labels[..., 1:], padding ignored; LlamaForCausalLM.forward behind train_fp8.py:278), the autograd Function must
hand the kernels contiguous 2-D operands and route dX / dW back in the input's shape, and CPU tensors must be refused (no CPU path)."""
import pytest
import torch
import torch.nn.functional as F

import llm_fp8_b200 as te
from llm_fp8_b200 import lm_head


def test_loss_module_shift_matches_hf(monkeypatch):
    seen = {}

    def fake(x, w, labels, ignore_index=-100):
        seen["shapes"] = (tuple(x.shape), tuple(w.shape), tuple(labels.shape))
        logits = (x.float() @ w.float().t())
        return F.cross_entropy(logits.reshape(-1, w.shape[0]), labels.reshape(-1), ignore_index=ignore_index)
    monkeypatch.setattr(lm_head, "lm_head_cross_entropy", fake)
    B, S, H, V = 2, 9, 16, 40
    head = te.LMHeadLoss(H, V, params_dtype=torch.float32)
    g = torch.Generator().manual_seed(0)
    hidden = torch.randn(B, S, H, generator=g)
    tokens = torch.randint(0, V, (B, S), generator=g)
    tokens[1, 5:] = -100                                   # padding at the end of the second sequence (labels = inputs with pads ignored)
    loss = head(hidden, tokens)
    logits = hidden @ head.weight.t()
    want = F.cross_entropy(logits[:, :-1].reshape(-1, V), tokens[:, 1:].reshape(-1), ignore_index=-100)
    assert torch.allclose(loss, want, rtol=1e-6, atol=1e-7)
    assert seen["shapes"] == ((B, S, H), (V, H), (B, S))    # no copy of the hidden states: the shift is done on the labels
    loss_noshift = head(hidden, tokens, shift=False)
    want = F.cross_entropy(logits.reshape(-1, V), tokens.reshape(-1), ignore_index=-100)
    assert torch.allclose(loss_noshift, want, rtol=1e-6, atol=1e-7)


def test_cpu_tensors_are_refused():
    x, w = torch.randn(8, 16).bfloat16(), torch.randn(32, 16).bfloat16()
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        te.lm_head_cross_entropy(x, w, torch.zeros(8, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        te.bf16_gemm(x, False, w, False, 8, 32, 16)
