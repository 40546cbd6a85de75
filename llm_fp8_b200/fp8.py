"""fp8_autocast and the delayed-scaling state (mirrors te.pytorch.fp8_autocast as used at te_llama.py:76,79).

State layout: every fp8 tensor of every module owns one *slot* in a flat per-(recipe, direction,
device) pool: amax_history[L, CAP], scale[CAP], scale_inv[CAP], fmax[CAP].  Quantise kernels atomically max into
amax_history[0, slot]; one K4 launch (llmfp8_amax_scale_update) updates the whole pool after ONE all-reduce(MAX) of row 0
over the data-parallel group — once per direction per step instead of once per autocast exit:

  * leaving the outermost fp8_autocast (grad enabled) marks the forward pools *pending*;
  * the backward of a module marks its backward pool pending;
  * a pool is flushed (all-reduce + K4) lazily, right before a module whose slots are pending quantises again.

For a module that runs once per step this is exactly TE's schedule (the scale a tensor is quantised with at step t was
computed from amaxes of steps < t), with 2 collectives per step instead of 2 per decoder layer.
"""
from __future__ import annotations

import os
import threading
from contextlib import contextmanager
from typing import Dict, List, Optional

import torch

from . import _lib
from .recipe import DelayedScaling, Format, MXFP8BlockScaling, Recipe, get_default_fp8_recipe

_POOL_CAP = 4096
# LLMFP8_NVLS_AMAX=1 (opt-in; not yet run on hardware): the amax history of every pool lives in symmetric, multicast-mapped memory and
# flush() is ONE launch — in-switch MAX over the data-parallel ranks + the scale update (csrc/nvls_reduce.cu: nvls_amax_scale_update_kernel)
# — instead of an NCCL all-reduce followed by K4.  Pools are created in the same order on every rank (first use in forward), which is
# what the collective set-up needs.
NVLS_AMAX = os.environ.get("LLMFP8_NVLS_AMAX", "0") == "1"


class ScalePool:
    """Flat delayed-scaling state for one (recipe, direction, device)."""

    def __init__(self, recipe: DelayedScaling, forward: bool, device: torch.device):
        self.recipe, self.forward, self.device = recipe, forward, device
        L = recipe.amax_history_len
        self.amax_history = torch.zeros(L, _POOL_CAP, dtype=torch.float32, device=device)
        self.scale = torch.ones(_POOL_CAP, dtype=torch.float32, device=device)
        self.scale_inv = torch.ones(_POOL_CAP, dtype=torch.float32, device=device)
        fmax = recipe.fp8_format.value.max_fwd if forward else recipe.fp8_format.value.max_bwd
        self.fmax = torch.full((_POOL_CAP,), fmax, dtype=torch.float32, device=device)
        self.used = 0
        self.pending = False          # amaxes recorded and their region closed, update not yet applied
        self.epoch = 0                # number of flushes
        self.group = None
        self._free = {}               # width -> released slot ranges
        self._nvls = None             # (multicast address of row 0, device array of signal-pad pointers, pad bytes, rank, world, handle, flat tensor)
        if NVLS_AMAX and device.type == "cuda" and recipe.reduce_amax and torch.distributed.is_available() \
                and torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
            self._setup_nvls(L)

    def _setup_nvls(self, L: int):
        """collective over the default group: re-homes amax_history in symmetric memory"""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        group = dist.group.WORLD
        try:
            symm.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass
        flat = symm.empty(L * _POOL_CAP, dtype=torch.float32, device=self.device)
        hdl = symm.rendezvous(flat, group=group)
        if not hdl.multicast_ptr:
            raise RuntimeError("LLMFP8_NVLS_AMAX=1: no NVLink multicast mapping for symmetric memory on this system")
        flat.zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier(group=group)
        self.amax_history = flat.view(L, _POOL_CAP)
        self._nvls = (int(hdl.multicast_ptr), int(hdl.signal_pad_ptrs_dev), int(symm.get_signal_pad_size()), dist.get_rank(group), dist.get_world_size(group),
                      hdl, flat)

    def claim(self, n: int) -> int:
        free = self._free.get(n)
        if free:
            off = free.pop()
            self.amax_history[:, off:off + n].zero_()          # a recycled range starts like a fresh one: scale 1, empty history
            self.scale[off:off + n].fill_(1.0)
            self.scale_inv[off:off + n].fill_(1.0)
            return off
        if self.used + n > _POOL_CAP:
            raise RuntimeError(f"fp8 scale pool exhausted ({_POOL_CAP} slots)")
        off = self.used
        self.used += n
        return off

    def release(self, off: int, n: int):
        """A module that claimed [off, off + n) is gone (Fp8Meta.__del__): the range is handed to the next claim of the same width."""
        self._free.setdefault(n, []).append(off)

    def bind_group(self, group):
        """The fp8_group of the enclosing fp8_autocast: the amax MAX exchange of this pool runs over it.  One group per pool."""
        if group is None or group is self.group:
            return
        if self.group is not None:
            raise RuntimeError("fp8_autocast: two different fp8_group values for modules sharing one delayed-scaling recipe pool")
        if self._nvls is not None:
            raise RuntimeError("LLMFP8_NVLS_AMAX=1 exchanges amaxes over the default group only; fp8_group is not supported with it")
        self.group = group

    # device pointers of a slot
    def scale_ptr(self, slot: int) -> int:
        return self.scale.data_ptr() + 4 * slot

    def scale_inv_ptr(self, slot: int) -> int:
        return self.scale_inv.data_ptr() + 4 * slot

    def amax_ptr(self, slot: int) -> int:
        return self.amax_history.data_ptr() + 4 * slot      # row 0

    def flush(self):
        """all-reduce(MAX) of this step's amaxes over the data-parallel group, then K4 over the used slots."""
        if not self.pending or self.used == 0:
            self.pending = False
            return
        T = self.used
        # scale_inv is VERSIONED: every update writes a fresh tensor (K4 writes all CAP entries and reads none), so a forward that
        # captured `pool.scale_inv` keeps de-scaling its saved fp8 bytes with the factors they were cast with however many updates
        # happen before its backward runs (fwd, fwd, bwd, bwd; an eval forward in between; recompute) — what TE gets from cloning
        # scale_inv in forward, without a copy kernel.  The old version dies with the last graph that holds it.
        self.scale_inv = torch.empty_like(self.scale_inv)
        if self._nvls is not None and self.group in (None, torch.distributed.group.WORLD):
            mc, pads, pad_bytes, rank, world, _, _ = self._nvls
            L, cap = self.amax_history.shape
            stream = torch.cuda.current_stream(self.device).cuda_stream
            rc = _lib.load().llmfp8_nvls_amax_scale_update(self.amax_history.data_ptr(), mc, self.scale.data_ptr(), self.scale_inv.data_ptr(),
                                                           self.fmax.data_ptr(), L, cap, self.recipe.margin, pads, pad_bytes, rank, world, 0, stream)
            _lib.check(rc, "nvls_amax_scale_update")
            from .dp import NvlsWatch
            NvlsWatch.get(self.device).poll()          # a timed-out meeting leaves the scales un-updated: raise (one flush later at worst)
            self.pending = False
            self.epoch += 1
            return
        if self.recipe.reduce_amax and torch.distributed.is_available() and torch.distributed.is_initialized():
            group = self.group
            if torch.distributed.get_world_size(group) > 1:
                torch.distributed.all_reduce(self.amax_history[0, :T], op=torch.distributed.ReduceOp.MAX, group=group)
        amax_scale_update(self.amax_history, self.scale, self.scale_inv, self.fmax, T, self.recipe.margin)
        self.pending = False
        self.epoch += 1

    def state(self) -> dict:
        T = self.used
        return {"amax_history": self.amax_history[:, :T].clone(), "scale": self.scale[:T].clone(), "scale_inv": self.scale_inv[:T].clone()}


def amax_scale_update(hist: torch.Tensor, scale: torch.Tensor, scale_inv: torch.Tensor, fmax: torch.Tensor, T: int, margin: int):
    """K4 over a pool laid out [L, CAP] / [CAP].  The kernel indexes hist densely as [L, width], so the full CAP width is
    updated in one launch: unused slots hold amax 0 and keep scale 1 (the amax > 0 guard)."""
    L, cap = hist.shape
    stream = torch.cuda.current_stream(hist.device).cuda_stream
    rc = _lib.load().llmfp8_amax_scale_update(hist.data_ptr(), scale.data_ptr(), scale_inv.data_ptr(), fmax.data_ptr(), L, cap, margin, stream)
    _lib.check(rc, "amax_scale_update")


class _State(threading.local):
    def __init__(self):
        self.stack: List[tuple] = []     # (enabled, recipe, group)
        self.touched_fwd: List[ScalePool] = []


_state = _State()
_pools: Dict[tuple, ScalePool] = {}
_pools_lock = threading.Lock()


def get_pool(recipe: DelayedScaling, forward: bool, device: torch.device) -> ScalePool:
    key = (recipe.key(), forward, device.index if device.index is not None else torch.cuda.current_device())
    with _pools_lock:
        p = _pools.get(key)
        if p is None:
            p = _pools[key] = ScalePool(recipe, forward, device)
        return p


def all_pools() -> List[ScalePool]:
    return list(_pools.values())


def reset_fp8_state():
    """Drop every delayed-scaling pool (tests)."""
    with _pools_lock:
        _pools.clear()


def flush_all():
    """Apply every pending amax reduction + scale update now (e.g. before checkpointing the scaling state)."""
    for p in all_pools():
        p.flush()


def is_fp8_enabled() -> bool:
    return bool(_state.stack) and _state.stack[-1][0]


def get_fp8_recipe() -> Recipe:
    if _state.stack and _state.stack[-1][1] is not None:
        return _state.stack[-1][1]
    return get_default_fp8_recipe()


def get_fp8_group():
    return _state.stack[-1][2] if _state.stack else None


def note_forward_pool(pool: ScalePool):
    if pool not in _state.touched_fwd:
        _state.touched_fwd.append(pool)


@contextmanager
def fp8_autocast(enabled: bool = True, calibrating: bool = False, fp8_recipe: Optional[Recipe] = None, fp8_group=None, _graph: bool = False):
    """Context manager selecting the FP8 recipe for the modules called inside it.  Nestable; the innermost recipe wins
    (the reference nests its per-block autocasts inside accelerate's outer one: train_fp8.py:184-188)."""
    if enabled and not torch.cuda.is_available():
        raise RuntimeError("fp8_autocast(enabled=True) needs a CUDA device: llm_fp8_b200 has no CPU path")
    if calibrating:
        raise NotImplementedError("fp8_autocast(calibrating=True) (amax collection during a higher-precision run) is not on the reference's path "
                                  "(te_llama.py:76,79 never passes it) and is not implemented")
    _state.stack.append((enabled, fp8_recipe, fp8_group))
    try:
        yield
    finally:
        _state.stack.pop()
        if not _state.stack:
            if torch.is_grad_enabled():
                for p in _state.touched_fwd:
                    p.pending = True
            _state.touched_fwd = []


def is_fp8_available(return_reason: bool = False):
    ok = torch.cuda.is_available() and torch.cuda.get_device_capability()[0] >= 10
    reason = "" if ok else "llm_fp8_b200 needs an sm_100a (B200) device"
    return (ok, reason) if return_reason else ok


def is_mxfp8_available(return_reason: bool = False):
    return is_fp8_available(return_reason)
