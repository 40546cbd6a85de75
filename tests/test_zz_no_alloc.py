"""llmfp8_rmsnorm_bwd_noalloc (include/llmfp8_b200.h): K5 with its fp32 workspace in a static device array inside the library — for callers
that cannot allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
Called through the C ABI (ctypes) and compared with the existing entry points llmfp8_rmsnorm_bwd / llmfp8_rmsnorm_bwd_add, which take a
caller workspace: same kernel, same grid, same partial-sum layout, so dx and dgamma must be IDENTICAL bytes (a fortiori within the 2e-2
band tests/test_gpu_quantize.py::test_rmsnorm_bwd uses against the fp32 reference), on two calls in a row — the second call meets the
first call's contents in the shared array.  Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(M, H, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, H, generator=g).bfloat16().cuda()
    dy = (torch.randn(M, H, generator=g) * 0.1).bfloat16().cuda()
    dres = (torch.randn(M, H, generator=g) * 0.1).bfloat16().cuda()
    gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16().cuda()
    return x, dy, dres, gamma


def _existing(lib, dy, x, gamma, rs, dres):
    M, H = x.shape
    ws = torch.empty(lib.llmfp8_rmsnorm_bwd_ws_rows(M) * H, dtype=torch.float32, device=x.device)
    dx, dg = torch.empty_like(x), torch.empty(H, dtype=torch.bfloat16, device=x.device)
    st = torch.cuda.current_stream().cuda_stream
    if dres is None:
        rc = lib.llmfp8_rmsnorm_bwd(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), rs.data_ptr(), dx.data_ptr(), dg.data_ptr(), ws.data_ptr(), M, H, st)
    else:
        rc = lib.llmfp8_rmsnorm_bwd_add(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), rs.data_ptr(), dres.data_ptr(), dx.data_ptr(), dg.data_ptr(),
                                        ws.data_ptr(), M, H, st)
    assert rc == 0, lib.llmfp8_last_error()
    return dx, dg


def _noalloc(lib, dy, x, gamma, rs, dres, dx, dg):
    M, H = x.shape
    rc = lib.llmfp8_rmsnorm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), rs.data_ptr(), dres.data_ptr() if dres is not None else None,
                                        dx.data_ptr(), dg.data_ptr(), M, H, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.llmfp8_last_error()


def test_rmsnorm_bwd_noalloc_matches_existing_entry_point_twice_in_a_row():
    from llm_fp8_b200 import _lib, ops
    lib = _lib.load()
    assert lib.llmfp8_rmsnorm_bwd_noalloc_max_floats() == 592 * 4096
    # two DIFFERENT problems back to back (the second one smaller: stale partial rows of the first stay in the array behind it), then the
    # first one again — every call must reproduce the caller-workspace entry point byte for byte
    cases = [(6144, 4096, True, 1), (300, 2048, False, 2), (6144, 4096, True, 1), (33, 4096, False, 3)]
    for M, H, with_dres, seed in cases:
        x, dy, dres, gamma = _inputs(M, H, seed)
        rs = ops.rmsnorm_rsigma(x, 1e-5)
        want_dx, want_dg = _existing(lib, dy, x, gamma, rs, dres if with_dres else None)
        dx, dg = torch.empty_like(x), torch.empty(H, dtype=torch.bfloat16, device="cuda")      # outputs are the caller's, allocated BEFORE the call
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats().get("num_alloc_retries", 0), torch.cuda.memory_allocated()
        _noalloc(lib, dy, x, gamma, rs, dres if with_dres else None, dx, dg)
        torch.cuda.synchronize()
        assert (torch.cuda.memory_stats().get("num_alloc_retries", 0), torch.cuda.memory_allocated()) == before     # nothing went through torch's allocator
        assert torch.equal(dx, want_dx), f"dx differs at M={M} H={H}: max |diff| {(dx.float() - want_dx.float()).abs().max().item():.3e}"
        assert torch.equal(dg, want_dg), f"dgamma differs at M={M} H={H}: max |diff| {(dg.float() - want_dg.float()).abs().max().item():.3e}"
        tol = 2e-2 * want_dx.float().abs().max().item()
        assert (dx.float() - want_dx.float()).abs().max().item() <= tol


def test_rmsnorm_bwd_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    from llm_fp8_b200 import _lib
    lib = _lib.load()
    M, H = 6144, 8192                                        # 592 partial rows x 8192 = twice the array
    assert lib.llmfp8_rmsnorm_bwd_ws_rows(M) * H > lib.llmfp8_rmsnorm_bwd_noalloc_max_floats()
    t = torch.zeros(16, dtype=torch.bfloat16, device="cuda")    # never touched: the size check comes first
    rc = lib.llmfp8_rmsnorm_bwd_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), None, t.data_ptr(), t.data_ptr(), M, H,
                                        torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.llmfp8_last_error().decode()
    assert "scratch" in msg and "llmfp8_rmsnorm_bwd" in msg, msg
    # a small M at the same H fits (rows x H within the maximum) and runs
    M2 = 200
    x, dy, dres, gamma = _inputs(M2, H, 5)
    from llm_fp8_b200 import ops
    rs = ops.rmsnorm_rsigma(x, 1e-5)
    want_dx, want_dg = _existing(lib, dy, x, gamma, rs, None)
    dx, dg = torch.empty_like(x), torch.empty(H, dtype=torch.bfloat16, device="cuda")
    _noalloc(lib, dy, x, gamma, rs, None, dx, dg)
    torch.cuda.synchronize()
    assert torch.equal(dx, want_dx) and torch.equal(dg, want_dg)
