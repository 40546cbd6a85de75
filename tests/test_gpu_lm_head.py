"""lm_head + cross-entropy (reference: HF LlamaForCausalLM.lm_head + CrossEntropyLoss behind train_fp8.py:278)
and the bf16 tcgen05 GEMM under it, against plain PyTorch fp32 references: GEMM outputs within 1e-2 of max|ref| on bf16 (1e-5 on fp32
outputs), the loss within 1e-3 relative, gradients within 2e-2 of max|ref| (bf16 d(logits) feeding two bf16 GEMMs)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rnd(shape, seed, std=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * std).bfloat16()


def check(got, want, tol):
    got, want = got.float().cpu(), want.float().cpu()
    err = (got - want).abs().max().item() / (want.abs().max().item() + 1e-30)
    assert err <= tol, f"rel err {err:.3e}"
    return err


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 512), (384, 768, 4096), (200, 264, 160), (1000, 3000, 2064), (6144, 4096, 1024)])
def test_bf16_gemm_layouts(M, N, K, a_mn, b_mn):
    import llm_fp8_b200 as te
    if (a_mn and M % 8) or (b_mn and N % 8):
        pytest.skip("MN-major operands need a 16 B pitch")
    a, b = rnd((M, K), 1).cuda(), rnd((N, K), 2, 0.1).cuda()
    want = (a.double() @ b.double().t()).float()
    A = a.t().contiguous() if a_mn else a
    B = b.t().contiguous() if b_mn else b
    check(te.bf16_gemm(A, a_mn, B, b_mn, M, N, K), want, 1e-2)
    check(te.bf16_gemm(A, a_mn, B, b_mn, M, N, K, out_dtype=torch.float32), want, 1e-5)


def test_bf16_gemm_bias_accumulate():
    import llm_fp8_b200 as te
    M, N, K = 320, 520, 256
    a, b, bias = rnd((M, K), 3).cuda(), rnd((N, K), 4).cuda(), rnd((N,), 5).cuda()
    base = (a.double() @ b.double().t()).float()
    check(te.bf16_gemm(a, False, b, False, M, N, K, bias=bias), base + bias.float(), 1e-2)
    out = torch.full((M, N), 3.0, dtype=torch.float32, device="cuda")
    te.bf16_gemm(a, False, b, False, M, N, K, out=out, accumulate=True)
    check(out, base + 3.0, 1e-5)


def _ref_loss(x, w, labels, ignore_index=-100):
    logits = (x.float() @ w.float().t()).bfloat16().float()             # HF: lm_head output in bf16, upcast for the loss
    return F.cross_entropy(logits, labels, ignore_index=ignore_index), logits


@pytest.mark.parametrize("M,H,V", [(512, 256, 1000), (256, 512, 4096), (130, 128, 264), (1024, 1024, 128256)])
def test_lm_head_cross_entropy_matches_torch(M, H, V):
    import llm_fp8_b200 as te
    g = torch.Generator().manual_seed(M + V)
    x = rnd((M, H), 11).cuda().requires_grad_(True)
    w = rnd((V, H), 12, 0.05).cuda().requires_grad_(True)
    labels = torch.randint(0, V, (M,), generator=g)
    labels[::7] = -100                                                   # padding positions
    labels[1] = V - 1
    labels[2] = 0
    labels = labels.cuda()
    loss = te.lm_head_cross_entropy(x, w, labels)
    xr, wr = x.detach().clone().requires_grad_(True), w.detach().clone().requires_grad_(True)
    want, _ = _ref_loss(xr, wr, labels)
    assert abs(loss.item() - want.item()) <= 1e-3 * abs(want.item()), (loss.item(), want.item())
    loss.backward()
    want.backward()
    check(x.grad, xr.grad, 2e-2)
    check(w.grad, wr.grad, 2e-2)
    assert x.grad.dtype == torch.bfloat16 and w.grad.dtype == torch.bfloat16
    # ignored rows carry no gradient
    assert torch.count_nonzero(x.grad[::7]) == 0


def test_lm_head_loss_module_shifts_like_hf():
    import llm_fp8_b200 as te
    B, S, H, V = 2, 64, 128, 512
    head = te.LMHeadLoss(H, V, device="cuda")
    hidden = rnd((B, S, H), 21).cuda()
    tokens = torch.randint(0, V, (B, S), generator=torch.Generator().manual_seed(3)).cuda()
    loss = head(hidden, tokens)                                          # labels = inputs, as the reference's collator sets them
    logits = (hidden.float() @ head.weight.float().t()).bfloat16().float()
    want = F.cross_entropy(logits[:, :-1].reshape(-1, V), tokens[:, 1:].reshape(-1))
    assert abs(loss.item() - want.item()) <= 1e-3 * abs(want.item())


def test_lm_head_all_rows_ignored_gives_zero():
    import llm_fp8_b200 as te
    x, w = rnd((64, 128), 1).cuda().requires_grad_(True), rnd((256, 128), 2).cuda().requires_grad_(True)
    labels = torch.full((64,), -100, dtype=torch.int64, device="cuda")
    loss = te.lm_head_cross_entropy(x, w, labels)
    loss.backward()
    assert loss.item() == 0.0 and torch.count_nonzero(x.grad) == 0 and torch.count_nonzero(w.grad) == 0
