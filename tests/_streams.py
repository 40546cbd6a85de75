"""CUDA stream / event semantics as a happens-before model, for checking on CPU that side-stream scheduling in the product waits where
it must (tests/test_stream_order_cpu.py).  Test infrastructure only.

`fake_cuda(monkeypatch)` replaces torch.cuda.current_stream / Stream / stream / Event / current_device (and Tensor.record_stream) with
objects that carry vector clocks: every tracked launch on a stream is an event of that stream; wait_stream / wait_event merge what the
waited-on side had seen.  `Tracker.op(reads, writes)` records a launch on the current stream; `Tracker.check()` then asserts

  * read-after-write: the latest launch that wrote a buffer happens-before every launch that reads it;
  * write-after-read / write-after-write on the same buffer are ordered too (a buffer overwritten while an earlier reader may still run);
  * `Tracker.assert_joined(stream)`: the stream has (transitively) waited for every launch so far — what must hold wherever untracked
    work (autograd's own kernels, the optimizer) continues on that stream.

Arithmetic still runs eagerly on CPU tensors, so values are always "right" here: only the ordering is under test."""
import contextlib
import itertools

import torch


class FakeStream:
    _ids = itertools.count(1)

    def __init__(self, device=None, name=None, **_kw):
        self.id = next(FakeStream._ids)
        self.name = name or f"stream{self.id}"
        self.clock = {self.id: 0}            # what this stream's next launch is ordered after: stream id -> number of launches

    @property
    def cuda_stream(self):
        return self.id

    def snapshot(self):
        return dict(self.clock)

    def _merge(self, clock):
        for k, v in clock.items():
            if v > self.clock.get(k, 0):
                self.clock[k] = v

    def wait_stream(self, other):
        self._merge(other.clock)

    def wait_event(self, ev):
        if ev.clock is None:
            raise AssertionError("waiting on an event that was never recorded")
        self._merge(ev.clock)

    def record_event(self, ev=None):
        ev = ev or FakeEvent()
        ev.record(self)
        return ev

    def launch(self):
        """-> (stream id, index, clock) of a new launch on this stream"""
        self.clock[self.id] += 1
        return self.id, self.clock[self.id], dict(self.clock)

    def synchronize(self):
        pass


class FakeEvent:
    def __init__(self, *a, **kw):
        self.clock = None

    def record(self, stream=None):
        self.clock = (stream or _STATE["current"]).snapshot()

    def synchronize(self):
        pass


_STATE = {"current": None, "main": None}


class Tracker:
    def __init__(self):
        self.launches = []                   # (name, stream id, index, clock, reads, writes)
        self._keep = []                      # every storage seen stays alive (the CPU allocator must not hand its address to a new
                                             # tensor) — the storage, not the tensor: autograd adopts a gradient only if it is its sole owner

    @staticmethod
    def _keys(tensors):
        out = set()
        for t in tensors:
            if isinstance(t, torch.Tensor):
                if t.numel():
                    out.add(t.untyped_storage().data_ptr())
            elif t is not None:
                out.add(t)                   # any hashable names a finer-grained resource (one slot of a pool, ...)
        return out

    def op(self, name, reads, writes):
        reads, writes = list(reads), list(writes)
        self._keep += [t.untyped_storage() for t in reads + writes if isinstance(t, torch.Tensor)]
        sid, idx, clock = _STATE["current"].launch()
        self.launches.append((name, sid, idx, clock, self._keys(reads), self._keys(writes)))

    @staticmethod
    def _before(a, b):
        """launch a happens-before launch b"""
        _, sid, idx, _, _, _ = a
        return b[3].get(sid, 0) >= idx and a is not b

    def check(self):
        last_write, readers = {}, {}
        for l in self.launches:
            name, _, _, _, reads, writes = l
            for k in reads:
                w = last_write.get(k)
                if w is not None and not self._before(w, l):
                    raise AssertionError(f"{name} reads a buffer that {w[0]} (stream {w[1]}) may still be writing")
                readers.setdefault(k, []).append(l)
            for k in writes:
                w = last_write.get(k)
                if w is not None and not self._before(w, l):
                    raise AssertionError(f"{name} overwrites a buffer that {w[0]} (stream {w[1]}) may still be writing")
                for r in readers.get(k, []):
                    if r is not l and not self._before(r, l):
                        raise AssertionError(f"{name} overwrites a buffer that {r[0]} (stream {r[1]}) may still be reading")
                last_write[k] = l
                readers[k] = []
        return len(self.launches)

    def assert_joined(self, stream=None):
        stream = stream or _STATE["main"]
        for l in self.launches:
            if stream.clock.get(l[1], 0) < l[2]:
                raise AssertionError(f"{stream.name} continues without having waited for {l[0]} (stream {l[1]})")

    def streams_used(self):
        return {l[1] for l in self.launches}


@contextlib.contextmanager
def _stream_ctx(s):
    prev = _STATE["current"]
    _STATE["current"] = s if s is not None else prev
    try:
        yield
    finally:
        _STATE["current"] = prev


def fake_cuda(monkeypatch):
    """-> Tracker; torch.cuda's stream surface now runs on the model above, with one 'main' stream current."""
    from llm_fp8_b200 import functional as fn
    main = FakeStream(name="main")
    _STATE["current"] = _STATE["main"] = main
    monkeypatch.setattr(torch.cuda, "current_stream", lambda device=None: _STATE["current"])
    monkeypatch.setattr(torch.cuda, "Stream", FakeStream)
    monkeypatch.setattr(torch.cuda, "stream", _stream_ctx)
    monkeypatch.setattr(torch.cuda, "Event", FakeEvent)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.Tensor, "record_stream", lambda self, stream: None)
    monkeypatch.setattr(fn, "_streams_on", lambda device: True)
    monkeypatch.setattr(fn, "_side_streams", {})
    return Tracker()


def tensors_in(obj):
    """every tensor reachable from an op argument / result (tensors, ops.MXTensor / ops.QTensor, tuples)"""
    if isinstance(obj, torch.Tensor):
        yield obj
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            yield from tensors_in(o)
    elif hasattr(obj, "__dataclass_fields__"):
        for f in obj.__dataclass_fields__:
            yield from tensors_in(getattr(obj, f))
