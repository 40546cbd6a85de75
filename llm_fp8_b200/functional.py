"""Autograd functions of the FP8 linear path: every forward / backward below is a fixed sequence of C-ABI calls
(ops.py).  The sequences follow SURVEY.md §3.2:

  Linear           fwd: K2(x) K2(W) G1            bwd: K2b(dY) G2 G3
  LayerNormLinear  fwd: K1(x) K2(W) G1            bwd: K2b(dY) G2 G3 K5
  LayerNormMLP     fwd: K1 K2(W1) G1+b K3 K2(W2) G1+b
                   bwd: K2b(dY) G2(fc2) K3b G3(fc2) G2(fc1) G3(fc1) K5
and their MXFP8 counterparts (K1m/K2m quantisers, block-scaled GEMMs).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import ops
from .fp8 import ScalePool, get_fp8_group, get_pool, note_forward_pool
from .recipe import DelayedScaling, MXFP8BlockScaling, Recipe, bwd_fmt, fwd_fmt


class Fp8Meta:
    """Slot bookkeeping of one module inside the per-recipe pools (claimed lazily, per recipe key)."""

    def __init__(self, n_fwd: int, n_bwd: int):
        self.n_fwd, self.n_bwd = n_fwd, n_bwd
        self._slots = {}          # (recipe key, device index) -> dict

    def get(self, recipe: DelayedScaling, device: torch.device):
        key = (recipe.key(), device.index)
        s = self._slots.get(key)
        if s is None:
            pf, pb = get_pool(recipe, True, device), get_pool(recipe, False, device)
            s = self._slots[key] = {"pf": pf, "pb": pb, "f0": pf.claim(self.n_fwd), "b0": pb.claim(self.n_bwd), "ef": -1, "eb": -1}
        return s

    def __del__(self):
        try:
            for s in self._slots.values():
                s["pf"].release(s["f0"], self.n_fwd)
                s["pb"].release(s["b0"], self.n_bwd)
        except Exception:          # interpreter shutdown
            pass

    @staticmethod
    def begin_fwd(s, training: bool):
        pf: ScalePool = s["pf"]
        group = get_fp8_group()
        if group is not None:
            pf.bind_group(group)
            s["pb"].bind_group(group)
        if pf.pending and s["ef"] == pf.epoch:
            pf.flush()
        if training:
            s["ef"] = pf.epoch
            note_forward_pool(pf)

    @staticmethod
    def begin_bwd(s):
        pb: ScalePool = s["pb"]
        if pb.pending and s["eb"] == pb.epoch:
            pb.flush()
        s["eb"] = pb.epoch

    @staticmethod
    def end_bwd(s):
        s["pb"].pending = True


def _2d(x: torch.Tensor) -> torch.Tensor:
    return x.reshape(-1, x.shape[-1])


def _rows(x: torch.Tensor) -> torch.Tensor:
    """[..., C] -> [R, C] readable by the strided quantisers: a column slice of a wider matrix (the query slice of a fused qkv
    output, an attention core's strided output) stays a view — no copy; anything else is made contiguous."""
    x2 = _2d(x)
    return x2 if (x2.is_contiguous() or ops.rows_ok(x2)) else x2.contiguous()


def _weight_pt(wcache, first, key, weight, ff, pf: ScalePool, slot) -> "ops.QTensor":
    """Per-tensor fp8 copy of a weight.  TE's `is_first_microbatch` contract: None = cast on every
    call; True = cast and keep; False = reuse the kept copy (gradient accumulation: the weights do not change between the
    micro-batches of one optimizer step).  A kept copy carries a snapshot of the scale_inv it was cast with, because the pool's
    slot moves on at the next flush."""
    if wcache is not None and first is False and key in wcache:
        return wcache[key]
    wq = ops.quantize(weight, ff, pf, slot)
    if wcache is not None and first is not None:
        wq = ops.QTensor(wq.data, wq.fmt, pf.scale_inv, slot)        # this version of scale_inv is never rewritten (ScalePool.flush)
        wcache[key] = wq
    return wq


def _weight_mx(wcache, first, key, weight, need_col: bool) -> "ops.MXTensor":
    """MXFP8 copies (rowwise for fprop, colwise for dgrad) of a weight under the same `is_first_microbatch` contract."""
    if wcache is not None:
        hit = wcache.get(key)
        if hit is not None and (hit.col is not None or not need_col):
            if first is False:
                return hit
            # copies deposited by optim.FusedAdamW in the optimizer's own pass over this weight: valid for every forward until the weight
            # changes again (same storage, same version counter — any in-place torch update since would have bumped it)
            if hit.src is not None and hit.src == (weight.data_ptr(), weight._version):
                return hit
    wq = ops.mx_quantize(weight, True, need_col)
    if wcache is not None and first is not None:
        wcache[key] = wq
    return wq


# LLMFP8_WGRAD_STREAM (default 1): dgrad and wgrad of one linear are independent.  With the wgrad GEMM issued on a side stream its CTAs fill the SMs that the dgrad
# GEMM's last, partial wave leaves idle (an [M, 4096] output at 8B shapes is 5.2 waves of 256 x 256 tiles: 14 % of that launch) and
# run under the RMSNorm backward / dSwiGLU that follow on the main stream.  The main stream rejoins before backward returns, so
# autograd (and the gradient all-reduce hooks) see the weight gradient in stream order as before.
WGRAD_SIDE_STREAM = os.environ.get("LLMFP8_WGRAD_STREAM", "1") == "1"
_side_streams = {}


def _streams_on(device: torch.device) -> bool:
    """side-stream scheduling applies (tests/_streams.py replaces this together with torch.cuda's stream objects to check the ordering)"""
    return device.type == "cuda"


class _SideWgrad:
    def __init__(self, device: torch.device):
        self.on = WGRAD_SIDE_STREAM and _streams_on(device)
        self.outs = []
        if self.on:
            idx = device.index if device.index is not None else torch.cuda.current_device()
            self.side = _side_streams.get(idx)
            if self.side is None:
                self.side = _side_streams[idx] = torch.cuda.Stream(device=idx)
            self.main = torch.cuda.current_stream(idx)

    def run(self, fn):
        """fn() launches one wgrad GEMM and returns its output; its operands were produced on the main stream before this call."""
        if not self.on:
            return fn()
        self.side.wait_stream(self.main)
        with torch.cuda.stream(self.side):
            out = fn()
        self.outs.append(out)
        return out

    def join(self):
        if self.on and self.outs:
            self.main.wait_stream(self.side)
            for o in self.outs:
                o.record_stream(self.main)       # allocated from the side stream's pool, consumed on the main stream


# Where a wgrad GEMM writes: by default a fresh tensor.  A data-parallel host may register, per weight buffer (keyed by its address — for
# the split qkv weights that is the joined buffer), a home for the gradient inside a symmetric, multicast-mapped arena
# (dp.NvlsGradArena): the GEMM's epilogue then stores dW straight into the memory the in-switch all-reduce works on, no copy in between.
# Empty unless such an arena exists; a miss (weights moved since) just allocates as before.  Entries: address -> (home, parameters whose
# gradient lives there).  While one of them still holds a .grad (accumulation over micro-batches) the home IS that gradient and must
# not be overwritten: the GEMM then gets a fresh tensor and autograd adds it in.
_WGRAD_OUT = {}


def wgrad_out(key: int, N: int, K: int, dtype: torch.dtype) -> Optional[torch.Tensor]:
    entry = _WGRAD_OUT.get(key)
    if entry is None:
        return None
    home, params = entry
    if home.dtype != dtype or home.numel() != N * K or any(p.grad is not None for p in params):
        return None
    return home.view(N, K)          # a fresh alias: autograd keeps it as .grad without a copy (sole owner, contiguous)


def _slice_of(x: torch.Tensor, x2: torch.Tensor):
    """x2 ([M, K], rows spaced) is columns [off, off + K) of a contiguous [M, W] matrix that x is a view of -> (W, off), else None.
    A plain Linear fed such a slice (the attention stand-in's query columns of the fused qkv output, layer._QuerySlice) can then hand
    its input gradient back as the same slice of a zero-padded [M, W] buffer — written in place by the dgrad GEMM's TMA stores — and
    the slicing op's backward returns that buffer instead of concatenating dq with zeros."""
    b = x._base
    if x2.is_contiguous() or b is None or not b.is_contiguous() or b.dim() < 2:
        return None
    W = b.shape[-1]
    off = x.storage_offset() - b.storage_offset()
    if x2.stride(0) != W or not (0 <= off <= W - x2.shape[1]) or off % 8 or b.numel() != x2.shape[0] * W:
        return None
    return W, off


def _padded_dx(M: int, K: int, wide, dtype, device) -> torch.Tensor:
    W, off = wide
    buf = torch.empty((M, W), dtype=dtype, device=device)
    if off:
        buf[:, :off].zero_()
    if off + K < W:
        buf[:, off + K:].zero_()
    buf._llmfp8_pad = (off, K)
    return buf[:, off:off + K]


def _wgrad(dyq, xq, N, K, T, weight):
    """dW[N,K] = dY^T X over T tokens; both operands read MN-major from the forward/backward fp8 copies."""
    return ops.fp8_gemm(dyq, True, xq, True, N, K, T, out_dtype=weight.dtype)


# ====================================================================================== per-tensor (DelayedScaling)
class LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, meta: Fp8Meta, recipe: DelayedScaling, ln_weight, eps, is_grad_enabled, residual=None, passthrough=False,
                wcache=None, first=None):
        """ln_weight None: plain Linear; else LayerNormLinear (RMSNorm fused into the input cast).
        residual [.., N]: y = residual + x W^T (+b), added in the GEMM epilogue; its gradient is dy itself.
        passthrough: also return an alias of x, so that the gradient of a residual branch taken from it arrives here and is folded
        into the RMSNorm backward instead of costing an accumulation pass."""
        dev = x.device
        s = meta.get(recipe, dev)
        Fp8Meta.begin_fwd(s, is_grad_enabled)
        pf, f0 = s["pf"], s["f0"]
        ff = fwd_fmt(recipe.fp8_format)
        x2 = _2d(x).contiguous() if ln_weight is not None else _rows(x)
        M, K = x2.shape
        N = weight.shape[0]
        rsigma = None
        if ln_weight is not None:
            xq, rsigma = ops.rmsnorm_quant(x2, ln_weight, eps, ff, pf, f0 + 0)
        else:
            xq = ops.quantize(x2, ff, pf, f0 + 0)
        wq = _weight_pt(wcache, first, ("w", recipe.key()), weight, ff, pf, f0 + 1)
        y = ops.fp8_gemm(xq, False, wq, False, M, N, K, bias=bias, residual=_rows(residual) if residual is not None else None)
        if is_grad_enabled:
            ctx.save_for_backward(xq.data, wq.data, x2 if ln_weight is not None else None, ln_weight, rsigma)
            ctx.meta, ctx.recipe, ctx.slots = meta, recipe, s
            ctx.has_bias = bias is not None
            ctx.shape = x.shape
            ctx.dims = (M, N, K)
            ctx.wdtype = weight.dtype
            ctx.wkey = weight.data_ptr()
            ctx.needs = (x.requires_grad, weight.requires_grad)
            ctx.has_residual = residual is not None
            ctx.wide = _slice_of(x, x2) if ln_weight is None else None
            ctx.w_si = (wq.scale_inv, wq.scale_inv_off)
            ctx.si = pf.scale_inv             # the version the saved fp8 bytes were cast under (ScalePool.flush replaces, never rewrites, it)
        ctx.set_materialize_grads(False)
        y = y.reshape(*x.shape[:-1], N)
        return (y, x.view_as(x)) if passthrough else y

    @staticmethod
    def backward(ctx, dy, dres=None):
        if dy is None:                       # only the pass-through alias was used downstream
            return (dres,) + (None,) * 11
        xq_d, wq_d, x2, ln_weight, rsigma = ctx.saved_tensors
        s, recipe = ctx.slots, ctx.recipe
        M, N, K = ctx.dims
        Fp8Meta.begin_bwd(s)
        pf, pb, f0, b0 = s["pf"], s["pb"], s["f0"], s["b0"]
        ff, bf = fwd_fmt(recipe.fp8_format), bwd_fmt(recipe.fp8_format)
        xq = ops.QTensor(xq_d, ff, ctx.si, f0 + 0)
        wq = ops.QTensor(wq_d, ff, *ctx.w_si)
        dy2 = _2d(dy).contiguous()
        dbias = None
        if ctx.has_bias:
            dyq, dbias = ops.quantize_bgrad(dy2, bf, pb, b0 + 0)
        else:
            dyq = ops.quantize(dy2, bf, pb, b0 + 0)
        wg = _SideWgrad(dy2.device)
        dx = ops.fp8_gemm(dyq, False, wq, True, M, K, N,                   # dgrad: contraction over N, W read MN-major
                          out=_padded_dx(M, K, ctx.wide, dy2.dtype, dy2.device) if ctx.wide is not None and dres is None else None)
        dw = wg.run(lambda: ops.fp8_gemm(dyq, True, xq, True, N, K, M, out=wgrad_out(ctx.wkey, N, K, ctx.wdtype), out_dtype=ctx.wdtype))   # wgrad: contraction over tokens
        dgamma = None
        if ln_weight is not None:
            dx, dgamma = ops.rmsnorm_bwd(dx, x2, ln_weight, rsigma, add=_2d(dres).contiguous() if dres is not None else None)
        elif dres is not None:
            dx = dx + _2d(dres)
        wg.join()
        Fp8Meta.end_bwd(s)
        return dx.reshape(ctx.shape), dw, dbias, None, None, dgamma, None, None, (dy if ctx.has_residual else None), None, None, None


class LayerNormMLPFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ln_weight, w1, b1, w2, b2, meta: Fp8Meta, recipe: DelayedScaling, eps, is_grad_enabled, residual=False,
                wcache=None, first=None):
        """residual=True: returns x + MLP(x) — the add rides in the fc2 GEMM epilogue, its gradient in the RMSNorm backward."""
        dev = x.device
        s = meta.get(recipe, dev)
        Fp8Meta.begin_fwd(s, is_grad_enabled)
        pf, f0 = s["pf"], s["f0"]
        ff = fwd_fmt(recipe.fp8_format)
        x2 = _2d(x).contiguous()
        M, H = x2.shape
        I2 = w1.shape[0]
        I = I2 // 2
        lnq, rsigma = ops.rmsnorm_quant(x2, ln_weight, eps, ff, pf, f0 + 0)
        w1q = _weight_pt(wcache, first, ("w1", recipe.key()), w1, ff, pf, f0 + 1)
        fc1 = ops.fp8_gemm(lnq, False, w1q, False, M, I2, H, bias=b1)
        aq = ops.swiglu_quant(fc1, ff, pf, f0 + 3)
        w2q = _weight_pt(wcache, first, ("w2", recipe.key()), w2, ff, pf, f0 + 4)
        y = ops.fp8_gemm(aq, False, w2q, False, M, H, I, bias=b2, residual=x2 if residual else None)
        if is_grad_enabled:
            ctx.save_for_backward(x2, ln_weight, rsigma, lnq.data, w1q.data, fc1, aq.data, w2q.data)
            ctx.recipe, ctx.slots = recipe, s
            ctx.residual = residual
            ctx.w1_si, ctx.w2_si = (w1q.scale_inv, w1q.scale_inv_off), (w2q.scale_inv, w2q.scale_inv_off)
            ctx.si = pf.scale_inv
            ctx.has_bias = b1 is not None
            ctx.shape = x.shape
            ctx.dims = (M, H, I)
            ctx.wdtype = w1.dtype
            ctx.wkey = (w1.data_ptr(), w2.data_ptr())
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, ln_weight, rsigma, lnq_d, w1q_d, fc1, aq_d, w2q_d = ctx.saved_tensors
        s, recipe = ctx.slots, ctx.recipe
        M, H, I = ctx.dims
        Fp8Meta.begin_bwd(s)
        pf, pb, f0, b0 = s["pf"], s["pb"], s["f0"], s["b0"]
        ff, bf = fwd_fmt(recipe.fp8_format), bwd_fmt(recipe.fp8_format)
        lnq = ops.QTensor(lnq_d, ff, ctx.si, f0 + 0)
        w1q = ops.QTensor(w1q_d, ff, *ctx.w1_si)
        aq = ops.QTensor(aq_d, ff, ctx.si, f0 + 3)
        w2q = ops.QTensor(w2q_d, ff, *ctx.w2_si)
        dy2 = _2d(dy).contiguous()
        db2 = db1 = None
        if ctx.has_bias:
            dyq, db2 = ops.quantize_bgrad(dy2, bf, pb, b0 + 2)
        else:
            dyq = ops.quantize(dy2, bf, pb, b0 + 2)
        da = ops.fp8_gemm(dyq, False, w2q, True, M, I, H)                        # fc2 dgrad
        dfc1q, db1 = ops.dswiglu_quant(da, fc1, bf, pb, b0 + 0, want_dbias=ctx.has_bias)
        wg = _SideWgrad(dy2.device)
        dw2 = wg.run(lambda: ops.fp8_gemm(dyq, True, aq, True, H, I, M, out=wgrad_out(ctx.wkey[1], H, I, ctx.wdtype), out_dtype=ctx.wdtype))   # fc2 wgrad
        dln = ops.fp8_gemm(dfc1q, False, w1q, True, M, H, 2 * I)                 # fc1 dgrad
        dw1 = wg.run(lambda: ops.fp8_gemm(dfc1q, True, lnq, True, 2 * I, H, M, out=wgrad_out(ctx.wkey[0], 2 * I, H, ctx.wdtype), out_dtype=ctx.wdtype))   # fc1 wgrad
        dx, dgamma = ops.rmsnorm_bwd(dln, x2, ln_weight, rsigma, add=dy2 if ctx.residual else None)
        wg.join()
        Fp8Meta.end_bwd(s)
        return dx.reshape(ctx.shape), dgamma, dw1, db1, dw2, db2, None, None, None, None, None, None, None, None, None


# ====================================================================================== MXFP8
# The block-scaled GEMM at 256-column tiles has ONE accumulator in TMEM (the scale factors take the rest), so its epilogue is on the
# critical path and the residual loads lengthen it: measured at Llama-3.1-8B shapes, +15 us (proj) / +35 us (fc2) against ~20 us for a
# separate add (profiles/r01_microbench_llama8b_v4_gemm.jsonl).  The MXFP8 functions therefore add the residual with a plain in-place
# add after the GEMM unless LLMFP8_MX_EPILOGUE_RESIDUAL=1; the per-tensor GEMM (two accumulators, epilogue hidden) fuses it (+7..10 us).
MX_EPILOGUE_RESIDUAL = os.environ.get("LLMFP8_MX_EPILOGUE_RESIDUAL", "1") == "1"


# SwiGLU / dSwiGLU fused with the block quantiser (csrc/mx_act_quantize.cu): bit-identical by construction to the two-kernel sequences
# below, minus the bf16 intermediate's round trip through HBM.  Not yet run on hardware -> off unless LLMFP8_MX_FUSED_ACT=1.
MX_FUSED_ACT = os.environ.get("LLMFP8_MX_FUSED_ACT", "1") == "1"


def _mx_gemm_residual(a, sfa, b, sfb, M, N, K, bias, residual):
    """fprop block-scaled GEMM (+bias) (+residual [M, N] bf16, rows possibly spaced)."""
    if residual is None or MX_EPILOGUE_RESIDUAL:
        return ops.mxfp8_gemm(a, sfa, False, b, sfb, False, M, N, K, bias=bias, residual=residual)
    y = ops.mxfp8_gemm(a, sfa, False, b, sfb, False, M, N, K, bias=bias)
    return y.add_(residual)


def _mx_quantize_dy(dy2: torch.Tensor, want_bgrad: bool):
    """row + col MXFP8 copies of dY and (for a biased Linear) the bias gradient from the same pass over dY"""
    if want_bgrad:
        return ops.mx_quantize_bgrad(dy2)
    return ops.mx_quantize(dy2, True, True), None


class MXLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, ln_weight, eps, is_grad_enabled, residual=None, passthrough=False, wcache=None, first=None):
        """residual / passthrough / wcache, first: as LinearFn.forward."""
        x2 = _2d(x).contiguous() if ln_weight is not None else _rows(x)
        M, K = x2.shape
        N = weight.shape[0]
        rsigma = None
        if ln_weight is not None:
            rsigma = ops.rmsnorm_rsigma(x2, eps)
            xq = ops.mx_quantize(x2, True, is_grad_enabled, rsigma, ln_weight)
        else:
            xq = ops.mx_quantize(x2, True, is_grad_enabled)
        wq = _weight_mx(wcache, first, ("w", "mx"), weight, is_grad_enabled)
        y = _mx_gemm_residual(xq.row, xq.sf_row, wq.row, wq.sf_row, M, N, K, bias, _rows(residual) if residual is not None else None)
        if is_grad_enabled:
            ctx.save_for_backward(xq.col, xq.sf_col, wq.col, wq.sf_col, x2 if ln_weight is not None else None, ln_weight, rsigma)
            ctx.has_bias = bias is not None
            ctx.shape, ctx.dims, ctx.wdtype = x.shape, (M, N, K), weight.dtype
            ctx.wkey = weight.data_ptr()
            ctx.has_residual = residual is not None
            ctx.wide = _slice_of(x, x2) if ln_weight is None else None
        ctx.set_materialize_grads(False)
        y = y.reshape(*x.shape[:-1], N)
        return (y, x.view_as(x)) if passthrough else y

    @staticmethod
    def backward(ctx, dy, dres=None):
        if dy is None:                       # only the pass-through alias was used downstream
            return (dres,) + (None,) * 9
        x_col, x_sfc, w_col, w_sfc, x2, ln_weight, rsigma = ctx.saved_tensors
        M, N, K = ctx.dims
        dy2 = _2d(dy).contiguous()
        dyq, dbias = _mx_quantize_dy(dy2, ctx.has_bias)
        # dgrad: dX[M,K] = dY[M,N] W[N,K]: dY rowwise (blocks along N), W colwise (blocks along N), W read MN-major
        wg = _SideWgrad(dy2.device)
        dx = ops.mxfp8_gemm(dyq.row, dyq.sf_row, False, w_col, w_sfc, True, M, K, N,
                            out=_padded_dx(M, K, ctx.wide, dy2.dtype, dy2.device) if ctx.wide is not None and dres is None else None)
        # wgrad: dW[N,K] = dY^T X: both colwise (blocks along tokens), both MN-major
        dw = wg.run(lambda: ops.mxfp8_gemm(dyq.col, dyq.sf_col, True, x_col, x_sfc, True, N, K, M, out=wgrad_out(ctx.wkey, N, K, ctx.wdtype), out_dtype=ctx.wdtype))
        dgamma = None
        if ln_weight is not None:
            dx, dgamma = ops.rmsnorm_bwd(dx, x2, ln_weight, rsigma, add=_2d(dres).contiguous() if dres is not None else None)
        elif dres is not None:
            dx = dx + _2d(dres)
        wg.join()
        return dx.reshape(ctx.shape), dw, dbias, dgamma, None, None, (dy if ctx.has_residual else None), None, None, None


class MXLayerNormMLPFn(torch.autograd.Function):
    """MXFP8 LayerNormMLP.  SwiGLU / dSwiGLU run in our pinned-exp kernel with a bf16 result that the block quantiser then
    reads (one extra bf16 round trip compared with a fully fused activation+MX quantiser: DESIGN.md 'next')."""

    @staticmethod
    def forward(ctx, x, ln_weight, w1, b1, w2, b2, eps, is_grad_enabled, residual=False, wcache=None, first=None):
        x2 = _2d(x).contiguous()
        M, H = x2.shape
        I2 = w1.shape[0]
        I = I2 // 2
        g = is_grad_enabled
        rsigma = ops.rmsnorm_rsigma(x2, eps)
        lnq = ops.mx_quantize(x2, True, g, rsigma, ln_weight)
        w1q = _weight_mx(wcache, first, ("w1", "mx"), w1, g)
        fc1 = ops.mxfp8_gemm(lnq.row, lnq.sf_row, False, w1q.row, w1q.sf_row, False, M, I2, H, bias=b1)
        if MX_FUSED_ACT:
            aq = ops.mx_swiglu_quantize(fc1, True, g)
        else:
            a = ops.swiglu_bf16(fc1)
            aq = ops.mx_quantize(a, True, g)
        w2q = _weight_mx(wcache, first, ("w2", "mx"), w2, g)
        y = _mx_gemm_residual(aq.row, aq.sf_row, w2q.row, w2q.sf_row, M, H, I, b2, x2 if residual else None)
        if g:
            ctx.save_for_backward(x2, ln_weight, rsigma, lnq.col, lnq.sf_col, w1q.col, w1q.sf_col, fc1, aq.col, aq.sf_col, w2q.col, w2q.sf_col)
            ctx.has_bias = b1 is not None
            ctx.residual = residual
            ctx.shape, ctx.dims, ctx.wdtype = x.shape, (M, H, I), w1.dtype
            ctx.wkey = (w1.data_ptr(), w2.data_ptr())
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, ln_weight, rsigma, ln_col, ln_sfc, w1_col, w1_sfc, fc1, a_col, a_sfc, w2_col, w2_sfc = ctx.saved_tensors
        M, H, I = ctx.dims
        dy2 = _2d(dy).contiguous()
        dyq, db2 = _mx_quantize_dy(dy2, ctx.has_bias)
        da = ops.mxfp8_gemm(dyq.row, dyq.sf_row, False, w2_col, w2_sfc, True, M, I, H)                       # fc2 dgrad
        if MX_FUSED_ACT and I % 128 == 0:
            dfq, db1 = ops.mx_dswiglu_quantize(da, fc1, True, True, want_dbias=ctx.has_bias)
        else:
            dfc1, db1 = ops.dswiglu_bf16(da, fc1, want_dbias=ctx.has_bias)
            dfq = ops.mx_quantize(dfc1, True, True)
        wg = _SideWgrad(dy2.device)
        dw2 = wg.run(lambda: ops.mxfp8_gemm(dyq.col, dyq.sf_col, True, a_col, a_sfc, True, H, I, M, out=wgrad_out(ctx.wkey[1], H, I, ctx.wdtype), out_dtype=ctx.wdtype))      # fc2 wgrad
        dln = ops.mxfp8_gemm(dfq.row, dfq.sf_row, False, w1_col, w1_sfc, True, M, H, 2 * I)                  # fc1 dgrad
        dw1 = wg.run(lambda: ops.mxfp8_gemm(dfq.col, dfq.sf_col, True, ln_col, ln_sfc, True, 2 * I, H, M, out=wgrad_out(ctx.wkey[0], 2 * I, H, ctx.wdtype), out_dtype=ctx.wdtype))  # fc1 wgrad
        dx, dgamma = ops.rmsnorm_bwd(dln, x2, ln_weight, rsigma, add=dy2 if ctx.residual else None)
        wg.join()
        return dx.reshape(ctx.shape), dgamma, dw1, db1, dw2, db2, None, None, None, None, None
