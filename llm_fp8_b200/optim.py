"""Fused AdamW + MXFP8 requantise: `FusedAdamW` is torch.optim.AdamW (the reference's optimizer,
train_fp8.py:200-204: bf16 parameters, bf16 state) whose pass over the GEMM weights of this package's modules — Linear.weight, LayerNormLinear's
(joined query | key | value) weight, LayerNormMLP's fc1_weight / fc2_weight — runs in ONE kernel per weight that also writes the weight's MXFP8
copies (csrc/adamw_mx.cu).  The next forward under an MXFP8BlockScaling recipe finds them in the module's cache (functional._weight_mx checks
storage address and version counter) and skips its weight casts: at Llama-3.1-8B shapes 148 us per decoder layer, 3.5 % of the step.  Every other
parameter (norm weights, biases, embeddings, lm_head) is stepped by a plain torch.optim.AdamW with the same hyper-parameters.

    opt = FusedAdamW(model, lr=1e-4, weight_decay=0.01)      # `model`: any nn.Module containing llm_fp8_b200 modules
    loss.backward(); opt.step(); opt.zero_grad()

Update rule (fp32 arithmetic, one rounding to bf16 per tensor; torch's own bf16 path rounds after every elementwise op, so results agree to
bf16 precision, not bit for bit):   m = b1 m + (1 - b1) g;   v = b2 v + (1 - b2) g^2;   w = w (1 - lr wd) - lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from . import _lib, modules as _m
from .ops import MXTensor, _mx_out, _p, _stream


def _gemm_weights(model: torch.nn.Module) -> List[Tuple[torch.Tensor, List[torch.nn.Parameter], "_m._Fp8Module", tuple]]:
    """-> [(weight buffer [out, in], the Parameters living in it, owning module, cache key of its MX copies)]"""
    out = []
    for mod in model.modules():
        if isinstance(mod, _m.LayerNormMLP):
            out += [(mod.fc1_weight, [mod.fc1_weight], mod, ("w1", "mx")), (mod.fc2_weight, [mod.fc2_weight], mod, ("w2", "mx"))]
        elif isinstance(mod, _m.LayerNormLinear):
            if mod._split is None:
                out.append((mod.weight, [mod.weight], mod, ("w", "mx")))
            elif _m._JOIN_SPLIT_WEIGHTS:
                ws = [getattr(mod, n + "weight") for n in mod._split]
                mod._joined = _m._joined_rows(ws, getattr(mod, "_joined", None))
                out.append((mod._joined, ws, mod, ("w", "mx")))
        elif isinstance(mod, _m.Linear):
            out.append((mod.weight, [mod.weight], mod, ("w", "mx")))
    ok = []
    for w, ps, mod, key in out:
        if w.is_cuda and w.dtype == torch.bfloat16 and w.shape[0] % 32 == 0 and w.shape[1] % 32 == 0 and all(p.requires_grad for p in ps):
            ok.append((w, ps, mod, key))
    return ok


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, model: torch.nn.Module, lr: float = 1e-3, betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, emit_mx_copies: bool = True):
        self.fused = _gemm_weights(model)
        fused_ids = {id(p) for _, ps, _, _ in self.fused for p in ps}
        rest = [p for p in model.parameters() if p.requires_grad and id(p) not in fused_ids]
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        params = [p for _, ps, _, _ in self.fused for p in ps]
        super().__init__([{"params": params}] if params else [{"params": rest}], defaults)
        self.emit = emit_mx_copies
        self._rest = torch.optim.AdamW(rest, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay) if (rest and params) else None
        self._only_rest = not params
        self._state = {}          # id(weight buffer) -> dict(step, exp_avg, exp_avg_sq)

    def _joined_grad(self, w: torch.Tensor, ps: List[torch.nn.Parameter]) -> Optional[torch.Tensor]:
        """the gradient of a weight buffer as one [out, in] tensor: the Parameters' grads are its row slices when our backward produced them"""
        gs = [p.grad for p in ps]
        if any(g is None for g in gs):
            return None
        if len(gs) == 1:
            return gs[0].contiguous()
        addr, item = gs[0].data_ptr(), gs[0].element_size()
        for g in gs:
            if not g.is_contiguous() or g.data_ptr() != addr:
                return torch.cat(gs, dim=0)
            addr += g.numel() * item
        base = gs[0]._base
        if base is not None and base.is_contiguous() and base.shape == w.shape and base.data_ptr() == gs[0].data_ptr():
            return base
        return torch.cat(gs, dim=0)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        if self._only_rest:
            raise RuntimeError("FusedAdamW: the model holds no llm_fp8_b200 GEMM weights on a CUDA device; use torch.optim.AdamW")
        g0 = self.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]
        lib = _lib.load()
        for w, ps, mod, key in self.fused:
            g = self._joined_grad(w, ps)
            if g is None:
                continue
            if g.dtype != torch.bfloat16:
                g = g.bfloat16()
            st = self._state.get(id(w))
            if st is None:
                st = self._state[id(w)] = {"step": 0, "exp_avg": torch.zeros_like(w), "exp_avg_sq": torch.zeros_like(w)}
            st["step"] += 1
            R, C = w.shape
            out = _mx_out(R, C, True, True, w.device) if self.emit else MXTensor(R, C)
            rc = lib.llmfp8_adamw_mx_step(w.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), R, C, lr, b1, b2, eps, wd,
                                          1.0 - b1 ** st["step"], 1.0 - b2 ** st["step"], _p(out.row), _p(out.sf_row), _p(out.col), _p(out.sf_col), _stream(w))
            _lib.check(rc, "adamw_mx_step")
            if self.emit:
                out.src = (w.data_ptr(), w._version)      # the kernel wrote through the raw pointer: the version counter still vouches for these bytes
                mod._wcache[key] = out
        if self._rest is not None:
            for grp in self._rest.param_groups:
                grp.update(lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd)
            self._rest.step()
        return loss

    def zero_grad(self, set_to_none: bool = True):
        super().zero_grad(set_to_none)
        if self._rest is not None:
            self._rest.zero_grad(set_to_none)
