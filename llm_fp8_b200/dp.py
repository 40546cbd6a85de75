"""Data-parallel plumbing of the FP8 linear path: one process per GPU, tokens (M) sharded over ranks,
weights replicated.  Only two exchanges exist on this path, both NCCL over NVLink:

  * amax MAX all-reduce of the delayed-scaling pools — fp8.ScalePool.flush(), once per direction per step;
  * weight-gradient all-reduce (each rank's dW = dY_local^T X_local is a partial sum over its tokens), launched per
    parameter the moment autograd has produced it so it overlaps the rest of the backward pass.

Mirrors what the reference gets from `DDP(model, gradient_as_bucket_view=True)` (train_multi_gpu.py:355-378) and its
env-var rendezvous (train_multi_gpu.py:969-1006), without the DDP wrapper: no bucket copies — the wgrad GEMM's output
tensor is the all-reduce buffer.
"""
from __future__ import annotations

import bisect
import contextlib
import os
import weakref
from typing import Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def init_from_env(backend: Optional[str] = None):
    """RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT rendezvous (same contract as train_multi_gpu.py:972-976).
    Returns (rank, world_size, local_rank); a no-op single-process setup when WORLD_SIZE is absent or 1."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        import datetime
        timeout = datetime.timedelta(seconds=int(os.environ.get("LLMFP8_DIST_TIMEOUT_S", "300")))   # fail fast on a mismatched collective
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local), timeout=timeout)
        else:
            dist.init_process_group(backend, rank=rank, world_size=world, timeout=timeout)
    elif torch.cuda.is_available():
        torch.cuda.set_device(local)
    return rank, world, local


def shard_tokens(global_batch: int, rank: int, world: int) -> range:
    """Sequences [lo, hi) of the global batch that `rank` owns (contiguous, remainder spread over the first ranks)."""
    base, rem = divmod(global_batch, world)
    lo = rank * base + min(rank, rem)
    return range(lo, lo + base + (1 if rank < rem else 0))


class GradAllReducer:
    """Averages parameter gradients over the data-parallel group, overlapped with backward.

    Large gradients (>= `direct_numel` elements) are all-reduced in place, individually, as soon as their
    post-accumulate hook fires (asynchronously, on NCCL's stream); small ones (norm weights, biases) are packed into one
    flat buffer and reduced together in `finish()`.  Call `finish()` after `backward()`; it makes the current stream wait
    for every outstanding reduction.

    Gradient accumulation (more than one backward() per optimizer step; DDP.no_sync in the reference, train_multi_gpu.py:714-725):
    run every micro-batch but the last inside `no_sync()` — the hooks then leave the locally accumulated gradient alone — and the
    last one outside it, so the SUM over micro-batches is reduced exactly once.  A hook that fires for a gradient tensor which was
    already reduced (or whose reduce is still in flight) raises instead of summing the reduced value a second time; after
    `optimizer.zero_grad(set_to_none=False)` (same tensor, zeroed in place) call `new_step()` to say so."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group=None, direct_numel: int = 1 << 20, average: bool = True,
                 prescaled: Optional[bool] = None, bucket_bytes: Optional[int] = None, nvls: Optional["NvlsGradArena"] = None):
        """prescaled (default; LLMFP8_GRAD_REDUCE=avg turns it off): the caller multiplies its loss by `loss_scale` (= 1/world) and
        the gradients are all-reduced with SUM — the same mean as AVG, but SUM is the one reduction NVSwitch can do in the fabric:
        NCCL's NVLS algorithms have no AVG / pre-multiplied sum, so with AVG it falls back to a 32-channel ring.  Measured at 8 B200s
        (mxfp8, same box): SUM -> NVLS / 24 channels 170.7 ms per step, AVG -> Ring 175.0 ms.  DDP itself pre-divides
        (train_multi_gpu.py:355-378).  A caller that ignores `loss_scale` gets summed, not averaged, gradients."""
        # bucket_bytes > 0 (LLMFP8_GRAD_BUCKET_MB; 0 = one collective per gradient): large gradients are held back until that many
        # bytes are pending and then all-reduced as ONE coalesced group (ncclGroupStart / End: one fused launch over the gradients' own
        # memory, still no copies) — 6 collectives per decoder layer become 1-2, at the price of starting each a little later.
        if bucket_bytes is None:
            # default 512 MB: measured on 8 B200s (profiles/r02/scale8*_*): one collective per gradient 171.9 ms per step (6.45x one GPU, 35 ms
            # of reductions still outstanding when backward returns); groups of 128 / 256 / 512 / 1024 MB 154.5 / 154.8 / 152.1 / 152.5 ms
            # (7.0-7.2x, 9-12 ms outstanding)
            bucket_bytes = int(float(os.environ.get("LLMFP8_GRAD_BUCKET_MB", "512")) * (1 << 20))
        self.bucket_bytes = max(0, int(bucket_bytes))
        self._pending: List[torch.Tensor] = []
        self._pending_bytes = 0
        self.collectives = 0
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if prescaled is None:
            prescaled = os.environ.get("LLMFP8_GRAD_REDUCE", "sum").lower() != "avg"
        self.prescaled = bool(prescaled) and average and self.world > 1
        self.loss_scale = 1.0 / self.world if self.prescaled else 1.0
        self.average = average and not self.prescaled
        self.direct_numel = direct_numel
        self._work: List = []
        self._small: List[torch.nn.Parameter] = []
        self._handles = []
        self.bytes_reduced = 0
        # nvls: gradients that the wgrad GEMMs wrote into the symmetric arena are summed in the NVSwitch fabric by our own kernel
        # (csrc/nvls_reduce.cu) on a side stream instead of going through NCCL; everything else takes the paths below.
        self.nvls = nvls if self.world > 1 else None
        self.nvls_launches = 0
        self._nvls_stream = None
        self._nvls_busy = False
        self.sync = True                # False inside no_sync(): hooks do nothing
        self._reduced = {}              # id(grad tensor) -> weakref: reduced (or being reduced) since it was created
        if self.nvls is not None and not self.prescaled:
            raise RuntimeError("the in-switch gradient reduce is a plain SUM: it needs prescaled=True (loss scaled by 1/world)")
        if self.world > 1:
            for p in params:
                if p.requires_grad:
                    self._handles.append(p.register_post_accumulate_grad_hook(self._hook))

    def _op(self):
        # gloo has no AVG: sum, then scale in finish()
        return dist.ReduceOp.AVG if (self.average and dist.get_backend(self.group) == "nccl") else dist.ReduceOp.SUM

    @contextlib.contextmanager
    def no_sync(self):
        """Micro-batches whose gradients only accumulate locally (every one but the last of an optimizer step)."""
        prev, self.sync = self.sync, False
        try:
            yield
        finally:
            self.sync = prev

    def new_step(self):
        """Forget which gradient tensors were reduced (needed only when gradients are zeroed in place instead of dropped)."""
        self._reduced = {}

    def _claim(self, g: torch.Tensor):
        ref = self._reduced.get(id(g))
        if ref is not None and ref() is g:
            raise RuntimeError("GradAllReducer: this gradient tensor was already all-reduced (or its reduce is still in flight) and backward "
                               "accumulated into it again — wrap all micro-batches but the last in reducer.no_sync(), and drop the gradients "
                               "(zero_grad(set_to_none=True)) or call reducer.new_step() between optimizer steps")
        if len(self._reduced) > 4096:
            self._reduced = {k: r for k, r in self._reduced.items() if r() is not None}
        self._reduced[id(g)] = weakref.ref(g)

    def _hook(self, p: torch.nn.Parameter):
        g = p.grad
        if g is None or not self.sync:
            return
        self._claim(g)
        if self.nvls is not None and g.is_contiguous():
            where = self.nvls.locate(g)
            if where is not None:
                if self._nvls_stream is None:
                    self._nvls_stream = self.nvls.new_stream()
                self._nvls_stream.wait_stream(self.nvls.current_stream())      # the wgrad GEMM that wrote g is on the current stream
                self.nvls.reduce(where, g, self._nvls_stream)
                self._nvls_busy = True
                self.nvls_launches += 1
                self.collectives += 1
                self.bytes_reduced += g.numel() * g.element_size()
                return
        if g.numel() >= self.direct_numel and g.is_contiguous():
            self.bytes_reduced += g.numel() * g.element_size()
            if self.bucket_bytes > 0:
                self._pending.append(g)
                self._pending_bytes += g.numel() * g.element_size()
                if self._pending_bytes >= self.bucket_bytes:
                    self._flush_pending()
            else:
                self._work.append((dist.all_reduce(g, op=self._op(), group=self.group, async_op=True), [g]))
                self.collectives += 1
        else:
            self._small.append(p)

    def _flush_pending(self):
        if not self._pending:
            return
        grads, self._pending, self._pending_bytes = self._pending, [], 0
        with dist._coalescing_manager(group=self.group, async_ops=True) as cm:
            for g in grads:
                dist.all_reduce(g, op=self._op(), group=self.group)
        self._work.append((cm, grads))
        self.collectives += 1

    def finish(self):
        if self.world == 1:
            return
        post_scale = self.average and self._op() == dist.ReduceOp.SUM
        if self._nvls_busy:
            self.nvls.current_stream().wait_stream(self._nvls_stream)
            self._nvls_busy = False
            self.nvls.poll()     # raises (at the latest one step on) if a rank meeting gave up
        self._flush_pending()
        if self._small:
            # norm weights and biases (a few hundred small tensors): ONE flat buffer, filled and emptied with one fused multi-tensor copy
            # each (torch._foreach_copy_) instead of a copy kernel per tensor — on the critical path after the last wgrad
            grads = [p.grad for p in self._small]
            flat = torch.empty(sum(g.numel() for g in grads), dtype=torch.float32, device=grads[0].device)
            views, off = [], 0
            for g in grads:
                views.append(flat[off:off + g.numel()].view_as(g))
                off += g.numel()
            torch._foreach_copy_(views, grads)
            dist.all_reduce(flat, op=self._op(), group=self.group)
            if post_scale:
                flat /= self.world
            torch._foreach_copy_(grads, views)
            self.bytes_reduced += flat.numel() * 4
            self.collectives += 1
            self._small = []
        for w, grads in self._work:
            w.wait()
            if post_scale:
                for g in grads:
                    g /= self.world
        self._work = []

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []


# ------------------------------------------------------------------------------------------ in-switch gradient reduce (NVLS)
class NvlsWatch:
    """Once-per-step poll of the library's timed-out-meeting counter without synchronising: each poll() first looks at the value the
    PREVIOUS poll copied (its event has long completed), raises if it is non-zero, then queues the next copy on `stream`.  A failure
    therefore surfaces one step late at worst — and it is sticky in the kernels (csrc/nvls_reduce.cu), so nothing computed after it
    is trusted either.  One watch per device."""
    _by_device = {}

    @classmethod
    def get(cls, device: torch.device) -> "NvlsWatch":
        idx = device.index if device.index is not None else torch.cuda.current_device()
        w = cls._by_device.get(idx)
        if w is None:
            w = cls._by_device[idx] = cls(idx)
        return w

    def __init__(self, idx: int):
        from . import _lib
        self._lib, self._check = _lib.load(), _lib.check
        self.idx = idx
        self._host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._event = None

    def poll(self, stream=None):
        if self._event is not None and self._event.query():
            self._raise_if(int(self._host[0]) & 0xFFFFFFFF)
        stream = stream or torch.cuda.current_stream(self.idx)
        with torch.cuda.device(self.idx):
            self._check(self._lib.llmfp8_nvls_timeouts_async(self._host.data_ptr(), stream.cuda_stream), "nvls_timeouts_async")
            self._event = torch.cuda.Event()
            self._event.record(stream)

    def check_now(self):
        import ctypes as C
        n = C.c_uint(0)
        with torch.cuda.device(self.idx):
            self._check(self._lib.llmfp8_nvls_timeouts(C.byref(n)), "nvls_timeouts")
        self._raise_if(n.value)

    @staticmethod
    def _raise_if(n: int):
        if n:
            raise RuntimeError(f"in-switch (NVLS) exchange: {n} rank meetings timed out or were refused after a time-out — the gradients / scales "
                               "since then are invalid; re-zero the signal pads and call llmfp8_nvls_reset, or fall back to NCCL")


def gemm_weight_buffers(module: torch.nn.Module) -> List[Tuple[torch.Tensor, List[torch.nn.Parameter]]]:
    """Every [out, in] buffer that a wgrad GEMM of the path writes a gradient for, with the Parameters living in it: Linear.weight,
    LayerNormLinear.weight (or the joined buffer of its split query / key / value weights: te_llama.py:198-211), LayerNormMLP's
    fc1_weight and fc2_weight.  Call it after the module is on its device (the joined buffer is established here if needed)."""
    from . import modules as m
    out = []
    for mod in module.modules():
        if isinstance(mod, m.LayerNormMLP):
            out += [(mod.fc1_weight, [mod.fc1_weight]), (mod.fc2_weight, [mod.fc2_weight])]
        elif isinstance(mod, m.LayerNormLinear):
            if mod._split is None:
                out.append((mod.weight, [mod.weight]))
            elif m._JOIN_SPLIT_WEIGHTS:
                ws = [getattr(mod, n + "weight") for n in mod._split]
                mod._joined = m._joined_rows(ws, getattr(mod, "_joined", None))
                out.append((mod._joined, ws))
        elif isinstance(mod, m.Linear):
            out.append((mod.weight, [mod.weight]))
    return [(w, ps) for w, ps in out if all(p.requires_grad for p in ps)]


def plan_arena(sizes: Sequence[int], chunk_bytes: int, align: int = 256) -> Tuple[List[Tuple[int, int]], List[int]]:
    """Packs buffers of `sizes` bytes, in order, into chunks of at most chunk_bytes (a buffer larger than that gets a chunk of its
    own) -> ([(chunk index, byte offset)] per buffer, [chunk size in bytes]).  Offsets are multiples of `align` (>= the 16 bytes the
    reduce kernel's vectors need).  Deterministic in its inputs: every rank computes the same plan, which is what makes the arena
    symmetric."""
    place, chunks = [], []
    used = None
    for n in sizes:
        n = -(-int(n) // align) * align
        if used is None or (used > 0 and used + n > chunk_bytes):
            chunks.append(0)
            used = 0
        place.append((len(chunks) - 1, used))
        used += n
        chunks[-1] = used
    return place, chunks


class NvlsGradArena:
    """Symmetric, multicast-mapped home of the large weight gradients (one identical layout on every rank).

    The wgrad GEMMs store dW straight into it (functional.wgrad_out), autograd adopts those views as `.grad`, and
    GradAllReducer(nvls=arena) sums each one across ranks inside the NVSwitch fabric with llmfp8_nvls_allreduce_sum_bf16
    (csrc/nvls_reduce.cu) — `max_ctas` CTAs of 128 threads (default 32, LLMFP8_NVLS_CTAS; small enough to share an SM with a resident GEMM CTA, where NCCL's
    24-32 CTAs each take a whole one) — on a side stream, overlapped with
    the rest of backward.  torch.distributed._symmetric_memory does the plumbing only: the VMM allocation, the peer / multicast
    mappings and the zeroed signal pads; every byte is moved by our kernel.

    bf16 gradients only; one arena per model; collective constructor (same buffers, same order on all ranks).
    STATUS: not yet run on hardware (written after round 1's GPU budget was spent): opt-in, bench.py builds it with LLMFP8_NVLS_REDUCE=1."""

    def __init__(self, buffers: Sequence[Tuple[torch.Tensor, List[torch.nn.Parameter]]], group=None, chunk_bytes: Optional[int] = None,
                 max_ctas: Optional[int] = None, timeout_ms: int = 0):
        if not (dist.is_initialized() and torch.cuda.is_available()):
            raise RuntimeError("NvlsGradArena needs CUDA and an initialised NCCL process group")
        import torch.distributed._symmetric_memory as symm

        from . import _lib
        self._lib = _lib.load()
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        buffers = [(w, ps) for w, ps in buffers if w.dtype == torch.bfloat16 and w.is_cuda and w.numel() % 8 == 0]
        if not buffers:
            raise RuntimeError("NvlsGradArena: no bf16 CUDA weight buffers to host")
        self.device = buffers[0][0].device
        if chunk_bytes is None:
            chunk_bytes = int(float(os.environ.get("LLMFP8_NVLS_CHUNK_MB", "1024")) * (1 << 20))
        self.max_ctas = int(os.environ.get("LLMFP8_NVLS_CTAS", "32")) if max_ctas is None else int(max_ctas)
        self.timeout_ms = int(timeout_ms)
        place, chunks = plan_arena([w.numel() * 2 for w, _ in buffers], chunk_bytes)
        try:
            symm.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass                                       # newer torch enables it implicitly
        self.pad_bytes = int(symm.get_signal_pad_size())
        self._chunks = []                              # (base address, bytes, multicast address, device array of signal-pad pointers, tensor, handle)
        with torch.cuda.device(self.device):
            for nbytes in chunks:
                t = symm.empty(nbytes // 2, dtype=torch.bfloat16, device=self.device)
                hdl = symm.rendezvous(t, group=self.group)
                mc = int(hdl.multicast_ptr or 0)
                if mc == 0:
                    raise RuntimeError("NvlsGradArena: this system has no NVLink multicast (NVLS) mapping for symmetric memory")
                t.zero_()
                self._chunks.append((t.data_ptr(), nbytes, mc, int(hdl.signal_pad_ptrs_dev), t, hdl))
            torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        self._bases = [c[0] for c in self._chunks]
        order = sorted(range(len(self._chunks)), key=lambda i: self._bases[i])
        self._chunks = [self._chunks[i] for i in order]
        self._bases = [c[0] for c in self._chunks]
        remap = {old: new for new, old in enumerate(order)}
        self.homes = []
        from . import functional as fn
        for (w, ps), (ci, off) in zip(buffers, place):
            t = self._chunks[remap[ci]][4]
            home = t[off // 2: off // 2 + w.numel()]
            self.homes.append(home)
            fn._WGRAD_OUT[w.data_ptr()] = (home, list(ps))
        self._keys = [w.data_ptr() for w, _ in buffers]
        self.bytes = sum(chunks)

    # -- used by GradAllReducer
    def current_stream(self):
        return torch.cuda.current_stream(self.device)

    def new_stream(self):
        return torch.cuda.Stream(device=self.device)

    def locate(self, g: torch.Tensor) -> Optional[Tuple[int, int]]:
        """(chunk, byte offset) if g lies wholly inside the arena and the reduce kernel can take it, else None."""
        return locate_in_chunks(self._bases, [c[1] for c in self._chunks], g.data_ptr(), g.numel() * g.element_size()) \
            if (g.dtype == torch.bfloat16 and g.numel() % 8 == 0) else None

    def reduce(self, where: Tuple[int, int], g: torch.Tensor, stream) -> None:
        from . import _lib
        ci, off = where
        _, _, mc, pads, _, _ = self._chunks[ci]
        rc = self._lib.llmfp8_nvls_allreduce_sum_bf16(mc + off, g.numel(), pads, self.pad_bytes, self.rank, self.world, self.max_ctas, self.timeout_ms,
                                                      stream.cuda_stream)
        _lib.check(rc, "nvls_allreduce_sum_bf16")

    def poll(self) -> None:
        """Non-blocking, once per step (GradAllReducer.finish): raises if a rank meeting was seen to have timed out."""
        NvlsWatch.get(self.device).poll(self.current_stream())

    def check(self) -> None:
        """Raises if any reduce launch gave up waiting for its peers (synchronises the device)."""
        NvlsWatch.get(self.device).check_now()

    def release(self) -> None:
        from . import functional as fn
        for k in self._keys:
            fn._WGRAD_OUT.pop(k, None)
        self._keys = []


def locate_in_chunks(bases: Sequence[int], sizes: Sequence[int], addr: int, nbytes: int) -> Optional[Tuple[int, int]]:
    """bases ascending.  -> (chunk index, byte offset) of [addr, addr + nbytes) if one chunk contains it and the offset is 16-byte aligned."""
    i = bisect.bisect_right(bases, addr) - 1
    if i < 0 or addr + nbytes > bases[i] + sizes[i] or (addr - bases[i]) % 16:
        return None
    return i, addr - bases[i]


class CAbiComm:
    """The path's two collectives through the C ABI's own NCCL communicator (llmfp8_nccl_*: include/llmfp8_b200.h) instead of
    torch.distributed — what a host without PyTorch's process groups binds.  The 128-byte id is created on rank 0 and handed
    round either by the caller (`unique_id=`) or over an already initialised torch.distributed group (any backend).

    Collectives run on `stream` (default: the current stream) and are in place; like every C-ABI call they are asynchronous."""

    def __init__(self, rank: int, world: int, device: torch.device, unique_id: Optional[bytes] = None, group=None):
        import ctypes as C

        from . import _lib
        self._lib = _lib.load()
        self.rank, self.world, self.device = rank, world, torch.device(device)
        if unique_id is None:
            buf = C.create_string_buffer(128)
            if rank == 0:
                _lib.check(self._lib.llmfp8_nccl_unique_id(buf), "nccl_unique_id")
            if world > 1:
                if not dist.is_initialized():
                    raise RuntimeError("CAbiComm needs either unique_id= or an initialised torch.distributed group to hand the id round")
                box = [buf.raw if rank == 0 else None]
                dist.broadcast_object_list(box, src=0, group=group)
                unique_id = box[0]
            else:
                unique_id = buf.raw
        if len(unique_id) != 128:
            raise ValueError("NCCL unique id must be 128 bytes")
        self.unique_id = bytes(unique_id)
        handle = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self._lib.llmfp8_nccl_init(self.unique_id, rank, world, C.byref(handle)), "nccl_init")
        self._comm = handle

    def _stream(self, stream) -> int:
        return (stream or torch.cuda.current_stream(self.device)).cuda_stream

    def all_reduce_max_f32(self, t: torch.Tensor, stream=None):
        """amax pools: element-wise MAX of a flat fp32 buffer (fp8.ScalePool.flush)."""
        from . import _lib
        if t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda:
            raise RuntimeError("all_reduce_max_f32 needs a contiguous CUDA fp32 tensor")
        _lib.check(self._lib.llmfp8_nccl_allreduce_max_f32(self._comm, t.data_ptr(), t.numel(), self._stream(stream)), "nccl_allreduce_max_f32")
        return t

    def all_reduce_sum_bf16(self, t: torch.Tensor, stream=None):
        """weight gradients: SUM in place (scale the loss by 1/world for the mean)."""
        from . import _lib
        if t.dtype != torch.bfloat16 or not t.is_contiguous() or not t.is_cuda:
            raise RuntimeError("all_reduce_sum_bf16 needs a contiguous CUDA bf16 tensor")
        _lib.check(self._lib.llmfp8_nccl_allreduce_sum_bf16(self._comm, t.data_ptr(), t.numel(), self._stream(stream)), "nccl_allreduce_sum_bf16")
        return t

    def destroy(self):
        if self._comm is not None and self._comm.value:
            from . import _lib
            _lib.check(self._lib.llmfp8_nccl_destroy(self._comm), "nccl_destroy")
        self._comm = None
