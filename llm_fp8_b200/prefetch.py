"""Weight copies quantised one decoder layer ahead, on a side stream (opt-in: LLMFP8_WEIGHT_PREFETCH=1 in bench.py; unmeasured).

The reference's training loop never passes `is_first_microbatch` (train_fp8.py:276-291), so TE — and this path by default — casts
every weight to FP8 / MXFP8 inside the forward that uses it: four memory-bound launches per decoder layer (8B: 886 MB of traffic,
~170 us, ~4 % of the step) sitting between the GEMMs.  The weights of layer i+1 do not depend on anything layer i computes, while the
GEMMs of layer i leave SMs idle in their last, partial wave (an [M, 4096] output is 5.19 waves of 256 x 256 tiles).  The prefetcher
issues layer i+1's casts on a side stream when layer i starts; they run under layer i's GEMMs, and layer i+1 finds the copies in its
modules' `_wcache` — the mechanism `is_first_microbatch=False` already uses (functional._weight_pt / _weight_mx), so the consuming
path is the validated one.  Same kernels, same scale / amax slots, same bytes: results are unchanged.

Layers 0 and 1 of every step cast inline: delayed scaling updates its pools when the first layer of a step runs, and a cast issued
before that would use last step's scales.  A layer whose recipe is not known yet (first step) or whose prefetch was not issued casts
inline as well — a miss is always just the default behaviour.

A prefetched copy is only used if nothing can have changed the weights since it was cast: it must come from the same pass over the
layers (a pass starts whenever layer 0 runs) and the parameters' version counters must be where they were (optimizer steps are
in-place updates).  Copies issued by a re-run of a layer outside the normal order (activation recomputation in backward) therefore
never reach the next step; they are dropped.
"""
from __future__ import annotations

import functools
from typing import List, Sequence

import torch

from . import functional as fn
from .modules import _Fp8Module


class WeightPrefetcher:
    def __init__(self, layers: Sequence[torch.nn.Module]):
        self.layers = list(layers)
        self._mods: List[List[_Fp8Module]] = [[m for m in layer.modules() if isinstance(m, _Fp8Module)] for layer in self.layers]
        self._ready = {}                 # layer index -> CUDA event (or True on CPU) of an issued, not yet consumed prefetch
        self._side = {}                  # device index -> side stream
        self.issued = self.consumed = self.rejected = 0
        self._pass = 0
        self._handles = [layer.register_forward_pre_hook(functools.partial(self._pre, i)) for i, layer in enumerate(self.layers)]

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []

    # -- forward pre-hook of layer idx: take what was prefetched for it, then start the next layer's casts
    def _pre(self, idx, module, args):
        if idx == 0:
            self._pass += 1
        entry = self._ready.pop(idx, None)
        if entry is not None:
            ev, issued_in, versions = entry
            mods = self._mods[idx]
            if ev is not True:
                torch.cuda.current_stream(ev_device(mods)).wait_event(ev)      # also orders the release of rejected copies
            if issued_in == self._pass and versions == _versions(mods):
                for m in mods:
                    if m._wcache:
                        m._prefetched = True
                self.consumed += 1
            else:
                for m in mods:
                    m._wcache.clear()
                self.rejected += 1
        if idx >= 1 and idx + 1 < len(self.layers):
            self._issue(idx + 1)

    def _issue(self, j):
        mods = [m for m in self._mods[j] if m._last_recipe is not None]
        if not mods:
            return
        dev = ev_device(mods)
        need_col = torch.is_grad_enabled()
        if not fn._streams_on(dev):      # CPU wiring tests: no streams, same bookkeeping
            for m in mods:
                m.prefetch_weights(m._last_recipe, dev, need_col)
            self._ready[j] = (True, self._pass, _versions(self._mods[j]))
            self.issued += 1
            return
        main = torch.cuda.current_stream(dev)
        side = self._side.get(dev.index)
        if side is None:
            side = self._side[dev.index] = torch.cuda.Stream(device=dev)
        side.wait_stream(main)           # optimizer updates, this step's scale update and the allocator's reuse of freed blocks are all on main
        with torch.cuda.stream(side):
            for m in mods:
                m.prefetch_weights(m._last_recipe, dev, need_col)
            ev = torch.cuda.Event()
            ev.record(side)
        for m in mods:
            for t in m._wcache_tensors():
                t.record_stream(main)    # allocated from the side stream's pool, consumed (and freed) on the main stream
        self._ready[j] = (ev, self._pass, _versions(self._mods[j]))
        self.issued += 1


def _versions(mods):
    return [p._version for m in mods for p in m.parameters(recurse=False)]


def ev_device(mods) -> torch.device:
    return next(mods[0].parameters()).device
