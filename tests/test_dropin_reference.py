"""Drop-in check: the reference's OWN decoder-swap files (te_llama.py, te_llama_hybrid.py, te_llama_mxfp8.py, te_qwen.py) are imported
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


# ------------------------------------------------------------------------------------------------ te_qwen.py
def _tiny_qwen_config(ref):
    return ref.Qwen3Config(vocab_size=128, hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=512,
                           max_position_embeddings=64, rms_norm_eps=1e-6, num_hidden_layers=2, head_dim=64)


def test_reference_te_qwen_builds_on_the_shim(no_cuda_move):
    """te_qwen.py:76-106 builds the same module pair through signature filtering (`_filter_kwargs_by_sig`, :60-66): every keyword
    it passes must be a NAMED parameter of our constructors or it is silently dropped — num_gqa_groups in particular (:88-90)."""
    import llm_fp8_b200 as ours
    ref = _load("te_qwen")
    cfg = _tiny_qwen_config(ref)
    layer = ref.TEQwen3DecoderLayer(cfg)
    mha, mlp = layer.self_attention, layer.layernorm_mlp
    assert isinstance(mha, ours.MultiheadAttention) and isinstance(mlp, ours.LayerNormMLP)
    assert mha.num_gqa_groups == 2 and mha.qkv_format == "bshd" and mha.layernorm_qkv.eps == 1e-6 and mha.dropout == 0.0
    assert mha.layernorm_qkv.query_weight.shape == (256, 256) and mha.layernorm_qkv.key_weight.shape == (128, 256)
    assert mha.proj.bias is None and mlp.fc1_weight.shape == (1024, 256) and mlp.fc2_weight.shape == (256, 512)
    assert ref.attn_recipe.fp8_format is ours.Format.HYBRID and ref.mlp_recipe.fp8_format is ours.Format.E4M3   # te_qwen.py:72-73
    # rope cache without position ids (te_qwen.py:108-133): [S, 1, 1, head_dim] table from RotaryPositionEmbedding(head_dim, rope_theta=?)
    rope = layer._get_rope_slice(16, torch.device("cpu"), None)
    assert isinstance(rope, torch.Tensor) and rope.shape == (16, 1, 1, 64)
    # the forward call filters its keywords by OUR forward signature (te_qwen.py:270-271): the KV-cache keywords must not get through
    import inspect
    allowed = set(inspect.signature(mha.forward).parameters)
    assert {"attention_mask", "rotary_pos_emb"} <= allowed and not ({"past_key_value", "use_cache", "output_attentions"} & allowed)
    import transformer_engine as te
    assert te.pytorch.is_fp8_available() in (True, False)            # te_qwen.py:268


def test_reference_te_qwen_causal_lm_and_replace_params(no_cuda_move):
    """TEQwen3ForCausalLM.__new__ (te_qwen.py:298-302) + replace_params (:346-389): gate rows first, then up rows, in fc1_weight."""
    ref = _load("te_qwen")
    cfg = _tiny_qwen_config(ref)
    model = ref.TEQwen3ForCausalLM(cfg)
    assert all(type(l).__name__ == "TEQwen3DecoderLayer" for l in model.model.layers)
    torch.manual_seed(0)
    sd = ref.Qwen3ForCausalLM(cfg).state_dict()
    ref.replace_params(sd, model.state_dict(), cfg)
    l1 = model.model.layers[1]
    I = cfg.intermediate_size
    assert torch.equal(l1.layernorm_mlp.fc1_weight[:I], sd["model.layers.1.mlp.gate_proj.weight"])
    assert torch.equal(l1.layernorm_mlp.fc1_weight[I:], sd["model.layers.1.mlp.up_proj.weight"])
    assert torch.equal(l1.layernorm_mlp.fc2_weight, sd["model.layers.1.mlp.down_proj.weight"])
    assert torch.equal(l1.self_attention.layernorm_qkv.key_weight, sd["model.layers.1.self_attn.k_proj.weight"])
    assert torch.equal(l1.self_attention.proj.weight, sd["model.layers.1.self_attn.o_proj.weight"])
    assert torch.equal(l1.layernorm_mlp.layer_norm_weight, sd["model.layers.1.post_attention_layernorm.weight"])


def test_reference_te_qwen_forward_without_gpu_fails_loudly(no_cuda_move):
    """No CPU fallback behind the Qwen swap either: in training mode (fp8 enabled, te_qwen.py:268-269) a CPU forward raises."""
    ref = _load("te_qwen")
    layer = ref.TEQwen3DecoderLayer(_tiny_qwen_config(ref)).train()
    with pytest.raises(RuntimeError):
        layer(torch.randn(2, 16, 256, dtype=torch.bfloat16))
