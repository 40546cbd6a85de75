"""LayerNormLinear(parameters_split=...): the query / key / value weight Parameters the reference addresses by name
(te_llama.py:198-211) are row slices of one buffer that feeds a single GEMM — host-side tensor logic, checked on CPU."""
import copy

import pytest
import torch

import llm_fp8_b200 as te
from llm_fp8_b200 import modules


def _module(dtype=torch.bfloat16):
    torch.manual_seed(0)
    return te.LayerNormLinear(64, 64 + 2 * 32, eps=1e-5, bias=False, normalization="RMSNorm",
                              parameters_split={"query": 64, "key": 32, "value": 32}, params_dtype=dtype)


def _aliased(m, w):
    off = 0
    for n in ("query_weight", "key_weight", "value_weight"):
        p = getattr(m, n)
        assert p.data_ptr() == w.data_ptr() + off * w.shape[1] * w.element_size(), n
        off += p.shape[0]
    assert off == w.shape[0]


def test_joined_weight_aliases_the_parameters_and_keeps_their_names():
    m = _module()
    before = {n: p.detach().clone() for n, p in m.named_parameters()}
    w, b = m._weight_bias()
    assert b is None and w.shape == (128, 64) and w.requires_grad
    _aliased(m, w)
    assert set(dict(m.named_parameters())) == {"layer_norm_weight", "query_weight", "key_weight", "value_weight"}
    assert set(m.state_dict()) >= {"layer_norm_weight", "query_weight", "key_weight", "value_weight"}
    assert not any(k.startswith("_joined") for k in m.state_dict())
    for n, p in m.named_parameters():
        assert torch.equal(p, before[n])                       # gathering moved the storage, not the values
    assert torch.equal(w, torch.cat([before["query_weight"], before["key_weight"], before["value_weight"]]))
    j = m._joined
    m._weight_bias()
    assert m._joined is j                                      # nothing is gathered again while the aliasing holds


def test_gradient_reaches_the_split_parameters_as_row_slices():
    m = _module(torch.float32)
    r = torch.randn(128, 64)
    w, _ = m._weight_bias()
    (w * r).sum().backward()
    got = torch.cat([m.query_weight.grad, m.key_weight.grad, m.value_weight.grad])
    assert torch.equal(got, r)
    # same result as the concatenating form
    m2 = _module(torch.float32)
    w2 = torch.cat([m2.query_weight, m2.key_weight, m2.value_weight])
    (w2 * r).sum().backward()
    for n in ("query_weight", "key_weight", "value_weight"):
        assert torch.equal(getattr(m, n).grad, getattr(m2, n).grad)
    # a second backward accumulates
    w, _ = m._weight_bias()
    (w * r).sum().backward()
    assert torch.equal(m.key_weight.grad, 2 * r[64:96])


def test_in_place_updates_keep_the_aliasing_and_module_apply_restores_it():
    m = _module()
    w, _ = m._weight_bias()
    j = m._joined
    with torch.no_grad():                                       # optimizer step / replace_params (.data.copy_, te_llama.py:196-211)
        m.key_weight.add_(1.0)
        m.value_weight.data.copy_(torch.full((32, 64), 3.0))
    w, _ = m._weight_bias()
    assert m._joined is j
    assert torch.equal(w[64:96], m.key_weight) and bool((w[96:] == 3.0).all())
    sd = {k: v.clone() for k, v in m.state_dict().items() if isinstance(v, torch.Tensor)}
    sd["query_weight"].zero_()
    m.load_state_dict(sd, strict=False)
    w, _ = m._weight_bias()
    assert m._joined is j and bool((w[:64] == 0).all())
    m = m.to(torch.float32)                                     # Module._apply: every Parameter gets its own storage again
    w, _ = m._weight_bias()
    assert m._joined is not j and w.dtype == torch.float32
    _aliased(m, w)
    assert bool((w[:64] == 0).all()) and bool((w[96:] == 3.0).all())
    m2 = copy.deepcopy(m)
    w2, _ = m2._weight_bias()
    _aliased(m2, w2)
    assert torch.equal(w2, w) and w2.data_ptr() != w.data_ptr()
    with torch.no_grad():
        m2.query_weight.fill_(7.0)
    assert bool((m._weight_bias()[0][:64] == 0).all())          # the copy does not write through to the original


def test_switch_restores_the_concatenating_form(monkeypatch):
    monkeypatch.setattr(modules, "_JOIN_SPLIT_WEIGHTS", False)
    m = _module(torch.float32)
    w, _ = m._weight_bias()
    assert w.data_ptr() != m.query_weight.data_ptr()
    w.sum().backward()
    assert torch.equal(m.value_weight.grad, torch.ones(32, 64))


def test_optimizer_step_through_the_joined_weight():
    m = _module(torch.float32)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-2)
    x = torch.randn(16, 64)
    losses = []
    for _ in range(5):
        opt.zero_grad()
        w, _ = m._weight_bias()
        loss = (x @ w.t()).pow(2).mean()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    assert losses[-1] < losses[0]
    _aliased(m, m._weight_bias()[0])
