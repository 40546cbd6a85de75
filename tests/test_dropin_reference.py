"""Drop-in check: the reference's OWN decoder-swap files (te_llama.py, te_llama_hybrid.py, te_llama_mxfp8.py) are imported
UNMODIFIED through the in-tree `transformer_engine` shim and build their TELlamaDecoderLayer / TELlamaForCausalLM on it.

Runs where /root/reference exists (the authoring container); skipped on the GPU box, where
tests/test_gpu_modules.py::test_decoder_layer_swap_runs exercises the same structure on the device.
No compute is done here (no GPU): construction, parameter names, replace_params aliasing, recipe objects."""
import importlib.util
import os
import sys

import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")


def _load(name):
    spec = importlib.util.spec_from_file_location(f"_ref_{name}", os.path.join(REF, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture()
def no_cuda_move(monkeypatch):
    # te_llama.py:66 calls .cuda() on the RoPE table at construction; there is no GPU in this container.
    monkeypatch.setattr(torch.Tensor, "cuda", lambda self, *a, **k: self)


def _tiny_config():
    from transformers import LlamaConfig
    return LlamaConfig(hidden_size=256, intermediate_size=512, num_attention_heads=4, num_key_value_heads=2, num_hidden_layers=2,
                       vocab_size=128, max_position_embeddings=64, rms_norm_eps=1e-5, tie_word_embeddings=True)


@pytest.mark.parametrize("fname", ["te_llama", "te_llama_hybrid", "te_llama_mxfp8"])
def test_reference_file_imports_and_builds(fname, no_cuda_move):
    import transformer_engine as te
    assert te.__file__.startswith(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "the in-tree shim must be the one imported"
    ref = _load(fname)
    import llm_fp8_b200 as ours
    cfg = _tiny_config()
    layer = ref.TELlamaDecoderLayer(cfg)
    assert isinstance(layer.self_attention, ours.MultiheadAttention) and isinstance(layer.layernorm_mlp, ours.LayerNormMLP)
    names = set(dict(layer.named_parameters()))
    assert names == {"self_attention.layernorm_qkv.layer_norm_weight", "self_attention.layernorm_qkv.query_weight",
                     "self_attention.layernorm_qkv.key_weight", "self_attention.layernorm_qkv.value_weight", "self_attention.proj.weight",
                     "layernorm_mlp.layer_norm_weight", "layernorm_mlp.fc1_weight", "layernorm_mlp.fc1_bias", "layernorm_mlp.fc2_weight",
                     "layernorm_mlp.fc2_bias"}
    assert layer.layernorm_mlp.eps == 1e-5 and layer.self_attention.layernorm_qkv.eps == cfg.rms_norm_eps
    assert layer.te_rope_emb.shape == (cfg.max_position_embeddings, 1, 1, cfg.hidden_size // cfg.num_attention_heads)
    # recipe objects the file builds at module level
    if fname == "te_llama":
        assert ref.attn_recipe.fp8_format is ours.Format.HYBRID and ref.mlp_recipe.fp8_format is ours.Format.E4M3
        assert ref.attn_recipe.amax_history_len == 16 and ref.mlp_recipe.amax_compute_algo == "max"
    elif fname == "te_llama_hybrid":
        assert ref.recipe.fp8_format is ours.Format.HYBRID
    else:
        assert isinstance(ref.mxfp8_recipe, ours.MXFP8BlockScaling)


def test_reference_causal_lm_and_replace_params(no_cuda_move):
    """TELlamaForCausalLM.__new__ (te_llama.py:95-98) + replace_params (te_llama.py:181-239) on a random HF state dict."""
    ref = _load("te_llama")
    from transformers import LlamaForCausalLM
    cfg = _tiny_config()
    torch.manual_seed(0)
    hf = LlamaForCausalLM(cfg)
    model = ref.TELlamaForCausalLM(cfg)
    assert type(model.model.layers[0]).__name__ == "TELlamaDecoderLayer"
    te_sd = model.state_dict()
    ref.replace_params(hf.state_dict(), te_sd, cfg)
    model.load_state_dict(te_sd, strict=False)
    H, I = cfg.hidden_size, cfg.intermediate_size
    for i in range(cfg.num_hidden_layers):
        hl, tl = hf.model.layers[i], model.model.layers[i]
        assert torch.equal(tl.self_attention.layernorm_qkv.query_weight, hl.self_attn.q_proj.weight)
        assert torch.equal(tl.self_attention.layernorm_qkv.key_weight, hl.self_attn.k_proj.weight)
        assert torch.equal(tl.self_attention.layernorm_qkv.value_weight, hl.self_attn.v_proj.weight)
        assert torch.equal(tl.self_attention.proj.weight, hl.self_attn.o_proj.weight)
        assert torch.equal(tl.self_attention.layernorm_qkv.layer_norm_weight, hl.input_layernorm.weight)
        assert torch.equal(tl.layernorm_mlp.layer_norm_weight, hl.post_attention_layernorm.weight)
        assert torch.equal(tl.layernorm_mlp.fc1_weight[:I], hl.mlp.gate_proj.weight)      # gate rows first
        assert torch.equal(tl.layernorm_mlp.fc1_weight[I:], hl.mlp.up_proj.weight)
        assert torch.equal(tl.layernorm_mlp.fc2_weight, hl.mlp.down_proj.weight)
        assert tl.layernorm_mlp.fc1_bias.abs().sum() == 0 and tl.layernorm_mlp.fc2_bias.shape == (H,)


def test_forward_without_gpu_fails_loudly(no_cuda_move):
    """No CPU fallback: running the reference layer without a CUDA device must raise, not compute on the host."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ref = _load("te_llama")
    layer = ref.TELlamaDecoderLayer(_tiny_config())
    with pytest.raises(RuntimeError, match="CUDA|no CPU"):
        layer(torch.zeros(1, 8, 256))
