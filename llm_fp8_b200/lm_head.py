"""lm_head + cross-entropy loss: the bf16 `[M, H] x [V, H]^T` head and loss the reference runs outside TE on
every step (HF `LlamaForCausalLM.lm_head` + `CrossEntropyLoss` behind `model(**batch)`, train_fp8.py:278; accelerate leaves it a bf16
`nn.Linear`).  Forward never materialises the `[M, V]` logits: the GEMM's epilogue keeps per-row softmax statistics (csrc/gemm.cu, EPI 1)
and a small kernel folds them into the log-sum-exp and the mean loss (csrc/lm_head.cu).  Backward runs the GEMM again with an epilogue that
emits d(logits) in bf16 (EPI 2), then dX and dW as two more bf16 tcgen05 GEMMs reading the operands in place (MN-major).

Numerics follow HF: logits are rounded to bf16 (the lm_head output dtype) before the fp32 cross-entropy; mean over the rows whose label is
not `ignore_index`."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .ops import _req, _req_out, _stream


def bf16_gemm(a: torch.Tensor, a_mn: bool, b: torch.Tensor, b_mn: bool, M: int, N: int, K: int, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None, accumulate: bool = False, out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """D[M, N] = A B^T (+bias) (+D), bf16 operands, fp32 accumulation.  a: [M, K] (or [K, M] when a_mn), b: [N, K] (or [K, N] when b_mn)."""
    a, b = _req(a, torch.bfloat16), _req(b, torch.bfloat16)
    if out is None:
        if accumulate:
            raise RuntimeError("accumulate=True needs an output tensor")
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    _req_out(out)
    rc = _lib.load().llmfp8_bf16_gemm(a.data_ptr(), a.shape[-1], int(a_mn), b.data_ptr(), b.shape[-1], int(b_mn), out.data_ptr(), out.stride(0),
                                      int(out.dtype == torch.float32), M, N, K, bias.data_ptr() if bias is not None else None, int(accumulate), 0,
                                      _stream(out))
    _lib.check(rc, "bf16_gemm")
    return out


class LMHeadCrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, labels, ignore_index):
        lib = _lib.load()
        x2 = _req(x.reshape(-1, x.shape[-1]).contiguous(), torch.bfloat16)
        w = _req(weight, torch.bfloat16)
        lab = labels.reshape(-1).contiguous()
        if lab.dtype != torch.int64 or not lab.is_cuda:
            raise RuntimeError("labels must be a CUDA int64 tensor")
        M, H = x2.shape
        V = w.shape[0]
        if lab.numel() != M or w.shape[1] != H:
            raise RuntimeError(f"lm_head: shapes do not match (x {tuple(x.shape)}, weight {tuple(w.shape)}, labels {tuple(labels.shape)})")
        dev = x2.device
        npart = lib.llmfp8_lm_head_partials(V)
        partials = torch.empty((npart, M, 2), dtype=torch.float32, device=dev)
        label_logit = torch.zeros(M, dtype=torch.float32, device=dev)
        lse = torch.empty(M, dtype=torch.float32, device=dev)
        ws = torch.empty(lib.llmfp8_ce_reduce_ws_floats(M), dtype=torch.float32, device=dev)
        out = torch.empty(2, dtype=torch.float32, device=dev)
        st = _stream(x2)
        _lib.check(lib.llmfp8_lm_head_ce_stats(x2.data_ptr(), H, w.data_ptr(), H, M, V, H, lab.data_ptr(), partials.data_ptr(), label_logit.data_ptr(), st),
                   "lm_head_ce_stats")
        _lib.check(lib.llmfp8_ce_reduce(partials.data_ptr(), npart, M, label_logit.data_ptr(), lab.data_ptr(), ignore_index, lse.data_ptr(), ws.data_ptr(),
                                        out.data_ptr(), st), "ce_reduce")
        ctx.save_for_backward(x2, w, lab, lse, out)
        ctx.ignore_index = ignore_index
        ctx.x_shape = x.shape
        return out[0]

    @staticmethod
    def backward(ctx, gout):
        lib = _lib.load()
        x2, w, lab, lse, out = ctx.saved_tensors
        M, H = x2.shape
        V = w.shape[0]
        # upstream gradient / number of counted rows, as a device scalar (no host read)
        gscale = (gout.float().reshape(1) / out[1:2].clamp_min(1.0)).contiguous()
        dlogits = torch.empty((M, V), dtype=torch.bfloat16, device=x2.device)
        _lib.check(lib.llmfp8_lm_head_dlogits(x2.data_ptr(), H, w.data_ptr(), H, M, V, H, lab.data_ptr(), ctx.ignore_index, lse.data_ptr(), gscale.data_ptr(),
                                              dlogits.data_ptr(), V, _stream(x2)), "lm_head_dlogits")
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = bf16_gemm(dlogits, False, w, True, M, H, V).reshape(ctx.x_shape)        # dX = dlogits W: contraction over V, W read MN-major
        if ctx.needs_input_grad[1]:
            dw = bf16_gemm(dlogits, True, x2, True, V, H, M, out_dtype=w.dtype)           # dW = dlogits^T X: contraction over tokens
        return dx, dw, None, None


def lm_head_cross_entropy(x: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """mean over counted rows of CE(bf16(x @ weight^T).float(), labels) — x [..., H] bf16, weight [V, H] bf16, labels [...] int64."""
    return LMHeadCrossEntropyFn.apply(x, weight, labels, ignore_index)


class LMHeadLoss(nn.Module):
    """`lm_head` weight [vocab, hidden] (bf16, no bias — as HF's) + the loss of HF `LlamaForCausalLM.forward(labels=...)`: with `shift=True`
    position t predicts token t + 1 (logits[..., :-1, :] against labels[..., 1:])."""

    def __init__(self, hidden_size: int, vocab_size: int, params_dtype: torch.dtype = torch.bfloat16, device=None, ignore_index: int = -100):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(vocab_size, hidden_size, dtype=params_dtype, device=device).normal_(0.0, 0.02))
        self.ignore_index = ignore_index

    def forward(self, hidden: torch.Tensor, labels: torch.Tensor, shift: bool = True) -> torch.Tensor:
        if shift:
            # same rows, labels moved one position left and the last position of every sequence ignored: no copy of the hidden states
            lab = torch.full_like(labels, self.ignore_index)
            lab[..., :-1] = labels[..., 1:]
            labels = lab
        return lm_head_cross_entropy(hidden, self.weight, labels, self.ignore_index)
