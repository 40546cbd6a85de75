"""Test helper: HF LlamaForCausalLM whose decoder layers are swapped for the FP8 layer structure the reference builds
(te_llama.py:41-82 and its hybrid / mxfp8 variants), with the HF weights mapped the way replace_params does
(te_llama.py:181-239).  Written against the `transformer_engine` shim so it reads like the reference's own swap; the
reference's files themselves are exercised unchanged in tests/test_dropin_reference.py (they do not exist on the GPU box)."""
import torch


def _recipes(scenario):
    from transformer_engine.common.recipe import DelayedScaling, Format, MXFP8BlockScaling
    if scenario == "default":           # te_llama.py:39-40
        return (DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max"),
                DelayedScaling(fp8_format=Format.E4M3, amax_history_len=16, amax_compute_algo="max"))
    if scenario == "hybrid":            # te_llama_hybrid.py:39
        r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max")
        return r, r
    r = MXFP8BlockScaling(fp8_format=Format.E4M3)     # te_llama_mxfp8.py:28-29
    return r, r


def make_layer_cls(scenario):
    import transformer_engine as te
    from transformer_engine.pytorch.attention import RotaryPositionEmbedding
    attn_recipe, mlp_recipe = _recipes(scenario)

    class SwappedDecoderLayer(torch.nn.Module):
        def __init__(self, config, *args, **kwargs):
            super().__init__()
            self.self_attention = te.pytorch.MultiheadAttention(
                hidden_size=config.hidden_size, num_attention_heads=config.num_attention_heads, bias=False,
                layernorm_epsilon=config.rms_norm_eps, attention_dropout=0.0, fuse_qkv_params=False, normalization="RMSNorm",
                num_gqa_groups=config.num_key_value_heads, qkv_format="bshd", input_layernorm=True)
            self.layernorm_mlp = te.pytorch.LayerNormMLP(hidden_size=config.hidden_size, ffn_hidden_size=config.intermediate_size,
                                                         normalization="RMSNorm", activation="swiglu")
            rope = RotaryPositionEmbedding(config.hidden_size // config.num_attention_heads)
            self.register_buffer("rope", rope(max_seq_len=config.max_position_embeddings), persistent=False)

        def forward(self, hidden_states, attention_mask=None, **kwargs):
            with te.pytorch.fp8_autocast(enabled=True, fp8_recipe=attn_recipe):
                a = self.self_attention(hidden_states, attention_mask=None, rotary_pos_emb=self.rope)
            hidden_states = hidden_states + a
            with te.pytorch.fp8_autocast(enabled=True, fp8_recipe=mlp_recipe):
                f = self.layernorm_mlp(hidden_states)
            return hidden_states + f

    return SwappedDecoderLayer


def build_swapped_causal_lm(tiny_config: dict, seed: int, scenario: str):
    """Same seeded HF init as tests/golden/make_golden.py:hf_bf16_losses, then the swap + weight map.  Returns a bf16 CPU model."""
    import transformers.models.llama.modeling_llama as ml
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(**tiny_config)
    torch.manual_seed(seed)
    hf = LlamaForCausalLM(cfg).to(torch.bfloat16)
    orig = ml.LlamaDecoderLayer
    old = torch.get_default_dtype()
    ml.LlamaDecoderLayer = make_layer_cls(scenario)
    torch.set_default_dtype(torch.bfloat16)          # te_llama.py:107
    try:
        model = LlamaForCausalLM(cfg)
    finally:
        ml.LlamaDecoderLayer = orig
        torch.set_default_dtype(old)
    model = model.to(torch.bfloat16)
    I = cfg.intermediate_size
    with torch.no_grad():
        model.model.embed_tokens.weight.copy_(hf.model.embed_tokens.weight)
        model.model.norm.weight.copy_(hf.model.norm.weight)
        model.lm_head.weight.copy_(hf.lm_head.weight)
        for tl, hl in zip(model.model.layers, hf.model.layers):
            qkv = tl.self_attention.layernorm_qkv
            qkv.layer_norm_weight.copy_(hl.input_layernorm.weight)
            qkv.query_weight.copy_(hl.self_attn.q_proj.weight)
            qkv.key_weight.copy_(hl.self_attn.k_proj.weight)
            qkv.value_weight.copy_(hl.self_attn.v_proj.weight)
            tl.self_attention.proj.weight.copy_(hl.self_attn.o_proj.weight)
            mlp = tl.layernorm_mlp
            mlp.layer_norm_weight.copy_(hl.post_attention_layernorm.weight)
            mlp.fc1_weight[:I].copy_(hl.mlp.gate_proj.weight)       # gate rows first (te_llama.py:225-233)
            mlp.fc1_weight[I:].copy_(hl.mlp.up_proj.weight)
            mlp.fc2_weight.copy_(hl.mlp.down_proj.weight)
            mlp.fc1_bias.zero_()
            mlp.fc2_bias.zero_()
    return model
