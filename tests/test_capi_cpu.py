"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/llmfp8_b200.h declares, the
header and the ctypes table agree, and the host-side contract errors fire without touching a GPU."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "llmfp8_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(llmfp8_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from llm_fp8_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in the header but not exported"
    assert sorted(_lib.EXPORTS) == syms, "ctypes prototype table and header disagree"
    lib.llmfp8_abi_version.restype = ctypes.c_int
    assert lib.llmfp8_abi_version() == 1


def test_pdl_build_exports_the_same_abi():
    """libllmfp8_b200_pdl.so (the -DLLMFP8_PDL build of the same sources, loaded with LLMFP8_PDL=1) is a drop-in for the default library."""
    path = os.path.join(ROOT, "llm_fp8_b200", "libllmfp8_b200_pdl.so")
    assert os.path.exists(path), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(path)
    for s in header_symbols():
        assert hasattr(lib, s), f"{s} missing from the PDL build"
    lib.llmfp8_abi_version.restype = ctypes.c_int
    assert lib.llmfp8_abi_version() == 1


def test_scale_bytes_matches_oracle():
    from llm_fp8_b200 import _lib, ops
    from oracle import fp8_oracle as fo
    lib = _lib.load()
    for rows, k in [(1, 32), (128, 128), (129, 160), (6144, 4096), (28672, 4096), (200, 96)]:
        assert lib.llmfp8_mx_scale_bytes(rows, k) == fo.sf_swizzled_size(rows, (k + 31) // 32) == ops.mx_scale_bytes(rows, k)


def test_no_cpu_fallback():
    import llm_fp8_b200 as te
    from llm_fp8_b200 import ops
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.quantize_raw(torch.zeros(4, 16, dtype=torch.bfloat16), 0, None, None)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="CUDA"):
            with te.fp8_autocast(enabled=True):
                pass


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "llm_fp8_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "liboracle" not in txt, f


def test_recipes_and_shim_surface():
    import transformer_engine as te
    from transformer_engine.common.recipe import DelayedScaling, Format, MXFP8BlockScaling
    from transformer_engine.pytorch.attention import RotaryPositionEmbedding
    r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max")
    assert r.margin == 0 and r.reduce_amax and Format.HYBRID.value.max_fwd == 448.0 and Format.HYBRID.value.max_bwd == 57344.0
    assert DelayedScaling().amax_history_len == 1024 and DelayedScaling().fp8_format is Format.HYBRID
    assert MXFP8BlockScaling(fp8_format=Format.E4M3).mxfp8()
    with pytest.raises(NotImplementedError):
        DelayedScaling(amax_compute_algo="most_recent")
    te.pytorch.TransformerLayer                     # te_llama.py:44 — the name must resolve
    for n in ("Linear", "LayerNormLinear", "LayerNormMLP", "MultiheadAttention", "fp8_autocast", "is_fp8_available"):
        assert hasattr(te.pytorch, n)
    f = RotaryPositionEmbedding(64)(max_seq_len=32)
    assert f.shape == (32, 1, 1, 64)


def test_parameter_names_and_aliasing():
    """replace_params (te_llama.py:181-239) writes through state_dict() entries with `.data[:] =`."""
    import transformer_engine as te
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        H, I, heads, kv = 256, 512, 4, 2
        attn = te.pytorch.MultiheadAttention(hidden_size=H, num_attention_heads=heads, bias=False, layernorm_epsilon=1e-5, attention_dropout=0.0,
                                             fuse_qkv_params=False, normalization="RMSNorm", num_gqa_groups=kv, qkv_format="bshd", input_layernorm=True)
        mlp = te.pytorch.LayerNormMLP(hidden_size=H, ffn_hidden_size=I, normalization="RMSNorm", activation="swiglu")
    finally:
        torch.set_default_dtype(old)
    sd = attn.state_dict()
    hd = H // heads
    want = {"layernorm_qkv.layer_norm_weight": (H,), "layernorm_qkv.query_weight": (H, H), "layernorm_qkv.key_weight": (kv * hd, H),
            "layernorm_qkv.value_weight": (kv * hd, H), "proj.weight": (H, H)}
    for k, shp in want.items():
        assert tuple(sd[k].shape) == shp and sd[k].dtype == torch.bfloat16
    assert not any(k.endswith("bias") for k in sd)
    sdm = mlp.state_dict()
    for k, shp in {"layer_norm_weight": (H,), "fc1_weight": (2 * I, H), "fc1_bias": (2 * I,), "fc2_weight": (H, I), "fc2_bias": (H,)}.items():
        assert tuple(sdm[k].shape) == shp
    assert (sdm["fc1_bias"] == 0).all() and mlp.eps == 1e-5
    sdm["fc1_weight"].data[:I] = 1.0          # gate rows
    sdm["fc1_weight"].data[I:] = 2.0          # up rows
    assert (mlp.fc1_weight[:I] == 1).all() and (mlp.fc1_weight[I:] == 2).all()
    sd["layernorm_qkv.query_weight"].data[:] = 3.0
    assert (attn.layernorm_qkv.query_weight == 3).all()
    missing, unexpected = attn.load_state_dict({k: v for k, v in sd.items() if not k.endswith("_extra_state")}, strict=False)
    assert not unexpected


def test_gemm_tuning_knobs_are_inert_without_opt_in():
    """llmfp8_gemm_debug / _force_bn / _force_cg change which kernel every thread's next launch gets: a process that did not set
    LLMFP8_ALLOW_TUNING=1 cannot be perturbed by them (checked in a fresh interpreter; this suite's conftest opts in)."""
    import subprocess
    import sys
    code = ("import os; os.environ.pop('LLMFP8_ALLOW_TUNING', None)\n"
            "from llm_fp8_b200 import _lib\n"
            "lib = _lib.load(); lib.llmfp8_gemm_debug(16)\n"
            "print(lib.llmfp8_last_error().decode())")
    env = {k: v for k, v in os.environ.items() if not k.startswith("LLMFP8_")}
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert "LLMFP8_ALLOW_TUNING" in out.stdout
