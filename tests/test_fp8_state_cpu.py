"""Host logic of fp8_autocast and the delayed-scaling pools (llm_fp8_b200/fp8.py, functional.Fp8Meta) on CPU: what te_llama.py:76,79
relies on from te.pytorch.fp8_autocast — nesting (the innermost recipe wins: the reference nests its per-block contexts inside
accelerate's outer one, train_fp8.py:184-188), thread-locality (autograd's backward thread), pools keyed by recipe, slot claims, the
"pending -> flush lazily before the next use" schedule (one amax exchange + one K4 launch per direction per step, and the scale used
at step t comes from amaxes of steps < t), and the amax MAX all-reduce inside ScalePool.flush with two gloo ranks.
The K4 kernel is replaced by the oracle's restatement (checker only); CUDA availability is faked — nothing here ships."""
import os
import socket
import threading

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import llm_fp8_b200 as te
from llm_fp8_b200 import fp8
from llm_fp8_b200 import functional as fn
from llm_fp8_b200.recipe import DelayedScaling, Format, MXFP8BlockScaling, bwd_fmt, fwd_fmt
from oracle import fp8_oracle as fo

CPU = torch.device("cpu", 0)


def f32div(a, b):
    return float(torch.tensor(a, dtype=torch.float32) / torch.tensor(b, dtype=torch.float32))


@pytest.fixture
def host_only(monkeypatch):
    """fp8.py with CUDA faked and K4 replaced by the oracle's update: the state machine runs on CPU tensors."""
    calls = []

    def k4(hist, scale, scale_inv, fmax, T, margin):
        calls.append(T)
        fo.amax_scale_update(hist, scale, scale_inv, fmax, margin)

    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(fp8, "amax_scale_update", k4)
    fp8.reset_fp8_state()
    yield calls
    fp8.reset_fp8_state()


def test_recipe_surface_matches_what_the_reference_passes():
    r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max")      # te_llama.py:39
    assert (r.margin, r.reduce_amax) == (0, True) and r.delayed() and not r.mxfp8()
    assert DelayedScaling().amax_history_len == 1024 and DelayedScaling().fp8_format is Format.HYBRID      # TE defaults
    assert Format.HYBRID.value == (448.0, 57344.0) and Format.E4M3.value == (448.0, 448.0) and Format.E5M2.value == (57344.0, 57344.0)
    assert (fwd_fmt(Format.HYBRID), bwd_fmt(Format.HYBRID)) == (0, 1) and (fwd_fmt(Format.E4M3), bwd_fmt(Format.E4M3)) == (0, 0)
    assert r.key() == DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max").key()
    assert r.key() != DelayedScaling(fp8_format=Format.E4M3, amax_history_len=16, amax_compute_algo="max").key()     # te_llama.py:40
    assert MXFP8BlockScaling(fp8_format=Format.E4M3).mxfp8()                                                        # te_llama_mxfp8.py:28-29
    with pytest.raises(NotImplementedError):
        DelayedScaling(amax_compute_algo="most_recent")
    with pytest.raises(NotImplementedError):
        MXFP8BlockScaling(fp8_format=Format.HYBRID)
    with pytest.raises(ValueError):
        DelayedScaling(amax_history_len=0)


def test_autocast_needs_a_gpu_when_enabled():
    if torch.cuda.is_available():
        pytest.skip("CPU-only contract")
    with pytest.raises(RuntimeError, match="no CPU path"):
        with te.fp8_autocast(enabled=True):
            pass
    with te.fp8_autocast(enabled=False):          # disabled contexts are fine anywhere
        assert not fp8.is_fp8_enabled()


def test_nesting_innermost_recipe_wins_and_exceptions_unwind(host_only):
    outer = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=16)
    inner = MXFP8BlockScaling()
    assert not fp8.is_fp8_enabled() and fp8.get_fp8_recipe().key() == DelayedScaling().key()
    with te.fp8_autocast(enabled=True, fp8_recipe=outer, fp8_group="g0"):
        assert fp8.is_fp8_enabled() and fp8.get_fp8_recipe() is outer and fp8.get_fp8_group() == "g0"
        with te.fp8_autocast(enabled=True, fp8_recipe=inner):
            assert fp8.get_fp8_recipe() is inner and fp8.get_fp8_group() is None
            with te.fp8_autocast(enabled=False):
                assert not fp8.is_fp8_enabled()
            assert fp8.is_fp8_enabled()
        assert fp8.get_fp8_recipe() is outer
        with pytest.raises(KeyError):
            with te.fp8_autocast(enabled=True, fp8_recipe=inner):
                raise KeyError("boom")
        assert fp8.get_fp8_recipe() is outer          # the failed context was popped
    assert not fp8.is_fp8_enabled()


def test_autocast_stack_is_thread_local(host_only):
    seen = {}

    def other():
        seen["enabled"] = fp8.is_fp8_enabled()
        with te.fp8_autocast(enabled=True, fp8_recipe=MXFP8BlockScaling()):
            seen["inner"] = fp8.get_fp8_recipe().mxfp8()

    r = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=4)
    with te.fp8_autocast(enabled=True, fp8_recipe=r):
        t = threading.Thread(target=other)
        t.start()
        t.join()
        assert fp8.get_fp8_recipe() is r               # the other thread's context did not leak in
    assert seen == {"enabled": False, "inner": True}


def test_pools_are_shared_per_recipe_and_slots_are_claimed_once(host_only):
    a = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16)
    b = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=16)
    pa, pa2, pb = fp8.get_pool(a, True, CPU), fp8.get_pool(DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16), True, CPU), fp8.get_pool(b, True, CPU)
    assert pa is pa2 and pa is not pb and fp8.get_pool(a, False, CPU) is not pa
    assert pa.amax_history.shape == (16, fp8._POOL_CAP) and float(pa.fmax[0]) == 448.0 and float(fp8.get_pool(a, False, CPU).fmax[0]) == 57344.0
    m1, m2 = fn.Fp8Meta(3, 2), fn.Fp8Meta(3, 2)
    s1, s2 = m1.get(a, CPU), m2.get(a, CPU)
    assert (s1["f0"], s1["b0"], s2["f0"], s2["b0"]) == (0, 0, 3, 2)
    assert m1.get(a, CPU) is s1                        # claimed once per (recipe, device)
    assert m1.get(b, CPU)["f0"] == 0                   # another recipe: another pool
    assert pa.scale_ptr(5) == pa.scale.data_ptr() + 20 and pa.amax_ptr(5) == pa.amax_history.data_ptr() + 20
    pa.used = fp8._POOL_CAP - 1
    with pytest.raises(RuntimeError, match="exhausted"):
        pa.claim(2)


def _fwd(meta, recipe, amax, training=True):
    """what a module's forward does with its slots: maybe flush, note the pool, quantise (here: record an amax in row 0)"""
    s = meta.get(recipe, CPU)
    fn.Fp8Meta.begin_fwd(s, training)
    used_scale = float(s["pf"].scale[s["f0"]])
    s["pf"].amax_history[0, s["f0"]] = max(float(s["pf"].amax_history[0, s["f0"]]), amax)
    return used_scale


def _bwd(meta, recipe, amax):
    s = meta.get(recipe, CPU)
    fn.Fp8Meta.begin_bwd(s)
    used_scale = float(s["pb"].scale[s["b0"]])
    s["pb"].amax_history[0, s["b0"]] = max(float(s["pb"].amax_history[0, s["b0"]]), amax)
    fn.Fp8Meta.end_bwd(s)
    return used_scale


def test_one_update_per_direction_per_step_and_scales_lag_by_one_step(host_only):
    calls = host_only
    r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=4)
    layers = [fn.Fp8Meta(3, 2) for _ in range(3)]
    amaxes = [2.0, 8.0, 4.0, 1.0]
    fwd_scales, bwd_scales = [], []
    for step, am in enumerate(amaxes):
        per_layer = []
        for m in layers:                                # the reference opens one autocast per block (te_llama.py:76,79)
            with te.fp8_autocast(enabled=True, fp8_recipe=r):
                per_layer.append(_fwd(m, r, am))
        assert len(set(per_layer)) == 1                 # every layer of a step sees the same generation of scales
        fwd_scales.append(per_layer[0])
        bwd_scales.append([_bwd(m, r, am * 10) for m in reversed(layers)][0])
        assert len(calls) == 2 * step                   # updates so far: one per direction per completed step
    assert fwd_scales == [1.0, f32div(448.0, 2.0), f32div(448.0, 8.0), f32div(448.0, 8.0)]          # step t uses max(amaxes of steps < t) over the history
    assert bwd_scales == [1.0, f32div(57344.0, 20.0), f32div(57344.0, 80.0), f32div(57344.0, 80.0)]
    assert calls == [9, 6] * 3                          # K4 covers every claimed slot of the pool in one launch
    fp8.flush_all()                                     # e.g. before checkpointing
    assert len(calls) == 8 and not any(p.pending for p in fp8.all_pools())


def test_no_grad_forward_does_not_schedule_an_update(host_only):
    calls = host_only
    r = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=4)
    m = fn.Fp8Meta(3, 2)
    for _ in range(3):
        with torch.no_grad():
            with te.fp8_autocast(enabled=True, fp8_recipe=r):
                _fwd(m, r, 3.0, training=False)
    assert calls == [] and not fp8.get_pool(r, True, CPU).pending
    with te.fp8_autocast(enabled=True, fp8_recipe=r):
        with te.fp8_autocast(enabled=True, fp8_recipe=r):
            _fwd(m, r, 3.0)
        assert not fp8.get_pool(r, True, CPU).pending   # only leaving the OUTERMOST context closes the region
    assert fp8.get_pool(r, True, CPU).pending


def test_module_extra_state_round_trips_the_scaling_state(host_only):
    r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=4)
    m = te.Linear(16, 16, bias=False, params_dtype=torch.bfloat16, device="cpu")
    s = m._meta.get(r, CPU)
    s["pf"].scale[s["f0"]:s["f0"] + 3] = torch.tensor([2.0, 3.0, 4.0])
    s["pf"].amax_history[1, s["f0"]] = 7.0
    sd = m.state_dict()
    assert any(k.endswith("_extra_state") for k in sd)
    fp8.reset_fp8_state()
    m2 = te.Linear(16, 16, bias=False, params_dtype=torch.bfloat16, device="cpu")
    m2.load_state_dict(sd)
    m2._apply_extra_state(r, CPU)
    s2 = m2._meta.get(r, CPU)
    assert s2["pf"].scale[s2["f0"]:s2["f0"] + 3].tolist() == [2.0, 3.0, 4.0] and float(s2["pf"].amax_history[1, s2["f0"]]) == 7.0
    assert torch.equal(m2.weight, m.weight)


# ------------------------------------------------------------------------------------------------ amax exchange, two gloo ranks
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _amax_worker(rank, world, port, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    calls = []

    def k4(hist, scale, scale_inv, fmax, T, margin):
        calls.append(T)
        fo.amax_scale_update(hist, scale, scale_inv, fmax, margin)

    fp8.amax_scale_update = k4
    r = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=4)
    local = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=4, reduce_amax=False)
    res = {}
    for name, rec in (("reduced", r), ("local", local)):
        pool = fp8.get_pool(rec, True, CPU)
        off = pool.claim(3)
        pool.amax_history[0, off:off + 3] = torch.tensor([1.0 + rank, 5.0 - rank, 0.0])
        pool.pending = True
        pool.flush()
        res[name] = pool.scale[off:off + 3].tolist()
    res["calls"] = calls
    torch.save(res, f"{out}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


def test_flush_exchanges_amax_so_every_rank_derives_the_same_scales(tmp_path):
    out = str(tmp_path / "r")
    mp.spawn(_amax_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert r0["reduced"] == r1["reduced"] == [f32div(448.0, 2.0), f32div(448.0, 5.0), 1.0]       # MAX over ranks; an unused slot keeps scale 1
    assert r0["local"] == [448.0, f32div(448.0, 5.0), 1.0] and r1["local"] == [f32div(448.0, 2.0), f32div(448.0, 4.0), 1.0]   # reduce_amax=False stays local
    assert r0["calls"] == r1["calls"] == [3, 3]




# ------------------------------------------------------------------------------------------------ saved fp8 copies keep THEIR scale_inv
class _ScaledOps:
    """Scale-aware stand-ins for the per-tensor C-ABI wrappers: the "fp8" copy is x * scale (fp32, no rounding), the amax lands in the
    pool's history row 0, the GEMM de-scales with whatever scale_inv tensor its operands carry.  With the right scale_inv every
    result is exact up to fp32 rounding; a stale or too-new one is off by the ratio of two scales."""

    @staticmethod
    def quantize(x, fmt, pool, slot):
        pool.amax_history[0, slot] = torch.maximum(pool.amax_history[0, slot], x.float().abs().max())
        return te.ops.QTensor(x.float() * pool.scale[slot], fmt, pool.scale_inv, slot)

    @staticmethod
    def quantize_bgrad(x, fmt, pool, slot):
        return _ScaledOps.quantize(x, fmt, pool, slot), x.float().sum(0).to(x.dtype)

    @staticmethod
    def fp8_gemm(a, a_mn, b, b_mn, M, N, K, bias=None, out=None, accumulate=False, out_dtype=torch.float32, max_ctas=0, residual=None):
        A = a.data.t() if a_mn else a.data
        B = b.data.t() if b_mn else b.data
        d = (A @ B.t()) * (a.scale_inv[a.scale_inv_off] * b.scale_inv[b.scale_inv_off])
        if bias is not None:
            d = d + bias.float()
        return d.to(torch.float32)


def test_backward_descales_saved_copies_with_the_scale_they_were_cast_with(host_only, monkeypatch):
    """fwd(x1), fwd(x2), bwd(2), bwd(1) on one Linear: the second forward applies the pending scale update, so the pool's scale_inv
    differs from the one x1 / W were cast under; backward 1 must still use the old one (TE clones scale_inv in forward for this;
    here every update writes a new scale_inv tensor and the graph keeps the version it saw)."""
    for name in ("quantize", "quantize_bgrad", "fp8_gemm"):
        monkeypatch.setattr(te.ops, name, getattr(_ScaledOps, name))
    monkeypatch.setattr(te.ops, "rows_ok", lambda t: True)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=4)
    g = torch.Generator().manual_seed(0)
    lin = te.Linear(32, 48, bias=True, params_dtype=torch.float32)
    x0 = torch.randn(8, 32, generator=g) * 3
    with te.fp8_autocast(enabled=True, fp8_recipe=r):        # step 0: records amaxes (scales still 1)
        lin(x0).sum().backward()
    lin.zero_grad()
    x1 = (torch.randn(8, 32, generator=g) * 10).requires_grad_(True)
    x2 = (torch.randn(8, 32, generator=g) * 100).requires_grad_(True)
    with te.fp8_autocast(enabled=True, fp8_recipe=r):
        y1 = lin(x1)                                       # applies step 0's update: scale = 448 / amax(x0)
    si_1 = fp8.all_pools()[0].scale_inv
    with te.fp8_autocast(enabled=True, fp8_recipe=r):
        y2 = lin(x2)                                       # applies y1's amaxes: a different scale, a NEW scale_inv tensor
    pf = next(p for p in fp8.all_pools() if p.forward)
    assert pf.scale_inv is not si_1 and not torch.equal(pf.scale_inv[:3], si_1[:3])
    dy1, dy2 = torch.randn(8, 48, generator=g), torch.randn(8, 48, generator=g)
    y2.backward(dy2)
    dw2 = lin.weight.grad.clone()
    lin.zero_grad()
    y1.backward(dy1)
    assert torch.allclose(lin.weight.grad, dy1.t() @ x1.detach(), rtol=1e-4, atol=1e-4)
    assert torch.allclose(x1.grad, dy1 @ lin.weight.detach(), rtol=1e-4, atol=1e-4)
    assert torch.allclose(dw2, dy2.t() @ x2.detach(), rtol=1e-4, atol=1e-3)
    assert torch.allclose(x2.grad, dy2 @ lin.weight.detach(), rtol=1e-4, atol=1e-4)


def test_released_slots_are_recycled_and_start_fresh(host_only):
    """building and dropping models in one process does not exhaust the 4096-slot pools (Fp8Meta.__del__ -> ScalePool.release)"""
    r = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=4)
    m = fn.Fp8Meta(3, 2)
    s = m.get(r, CPU)
    pf, f0 = s["pf"], s["f0"]
    pf.scale[f0:f0 + 3] = 7.0
    pf.amax_history[:, f0:f0 + 3] = 5.0
    used = pf.used
    del m, s
    import gc
    gc.collect()
    m2 = fn.Fp8Meta(3, 2)
    s2 = m2.get(r, CPU)
    assert s2["f0"] == f0 and pf.used == used
    assert torch.all(pf.scale[f0:f0 + 3] == 1) and torch.all(pf.amax_history[:, f0:f0 + 3] == 0)
    m3 = fn.Fp8Meta(6, 4)                      # another width: a fresh range
    assert m3.get(r, CPU)["f0"] == used


def test_fp8_group_binds_the_pool_and_conflicts_raise(host_only):
    r = DelayedScaling(fp8_format=Format.E4M3, amax_history_len=4)
    m = fn.Fp8Meta(3, 2)
    ga, gb = object(), object()
    with te.fp8_autocast(enabled=True, fp8_recipe=r, fp8_group=ga):
        s = m.get(r, CPU)
        fn.Fp8Meta.begin_fwd(s, True)
    assert s["pf"].group is ga and s["pb"].group is ga
    with te.fp8_autocast(enabled=True, fp8_recipe=r, fp8_group=gb):
        with pytest.raises(RuntimeError, match="two different fp8_group"):
            fn.Fp8Meta.begin_fwd(s, True)
    with pytest.raises(NotImplementedError):
        with te.fp8_autocast(enabled=True, fp8_recipe=r, calibrating=True):
            pass


def test_foreign_extra_state_is_ignored_with_a_warning():
    lin = te.Linear(16, 16, params_dtype=torch.float32)
    with pytest.warns(UserWarning, match="ignoring an _extra_state"):
        lin.set_extra_state(torch.arange(10, dtype=torch.uint8))       # what a TE-written checkpoint carries
    assert lin._pending_extra_state is None
    lin.set_extra_state(torch.empty(0, dtype=torch.uint8))
    lin.set_extra_state(None)
    lin.set_extra_state({"x": 1})
    assert lin._pending_extra_state == {"x": 1}
