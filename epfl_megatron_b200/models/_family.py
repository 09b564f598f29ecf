"""Architecture contracts of the named model families.

A family (Llama, Falcon) is a GPTModel plus a set of argument values that must hold (a wrong one
would silently build a different architecture than the checkpoint describes) and a set that is
merely unusual (warned about). Both are data here; ``check_family`` applies them.
"""
import warnings

from .enums import PositionEmbeddingType

# (argument, predicate on its value, what the family needs)
REQUIRED = {
    "Llama": [
        ("position_embedding_type", lambda v: v == PositionEmbeddingType.rotary, "rotary position embeddings"),
        ("use_post_ln", lambda v: not v, "pre-normalisation (no --use_post_ln)"),
        ("glu_activation", lambda v: v == "swiglu", "--glu_activation swiglu"),
        ("use_bias", lambda v: not v, "no biases"),
        ("parallel_attn", lambda v: not v, "sequential attention / MLP (no --parallel_attn)"),
        ("use_rms_norm", bool, "--use_rms_norm"),
        ("tie_embed_logits", lambda v: not v, "an untied LM head (--no_tie_embed_logits)"),
    ],
    "Falcon": [
        ("position_embedding_type", lambda v: v == PositionEmbeddingType.rotary, "rotary position embeddings"),
        ("num_attention_heads_kv", lambda v: isinstance(v, int), "an explicit --num_attention_heads_kv"),
        ("use_post_ln", lambda v: not v, "pre-normalisation (no --use_post_ln)"),
        ("glu_activation", lambda v: v is None, "a GeLU MLP (no --glu_activation)"),
        ("use_bias", lambda v: not v, "no biases"),
        ("parallel_attn", bool, "--parallel_attn"),
    ],
}

# (predicate on args, message)
UNUSUAL = {
    "Llama": [
        (lambda a: not a.use_flash_attn, "Llama trains fastest with the fused attention kernel (--use_flash_attn)"),
        (lambda a: not a.bf16, "Llama is meant to be trained in bf16"),
        (lambda a: a.hidden_dropout > 0.0 and not a.lima_dropout, "Llama is not intended to use dropout"),
        (lambda a: a.attention_dropout > 0.0, "Llama is not intended to use dropout"),
    ],
    "Falcon": [
        (lambda a: not a.parallel_layernorm, "Falcon uses parallel_layernorm, or are you running falcon-7b?"),
        (lambda a: not a.use_flash_attn, "Falcon trains fastest with the fused attention kernel (--use_flash_attn)"),
        (lambda a: a.bias_gelu_fusion, "Falcon has no MLP bias, so bias_gelu_fusion has no effect"),
        (lambda a: a.bias_dropout_fusion, "Falcon has no biases, so bias_dropout_fusion has no effect"),
        (lambda a: a.hidden_dropout > 0.0 and not a.lima_dropout, "Falcon is not intended to use dropout"),
    ],
}


def check_family(family: str, args) -> None:
    for name, ok, need in REQUIRED[family]:
        value = getattr(args, name)
        assert ok(value), f"{family} needs {need}; got {name}={value}"
    for odd, message in UNUSUAL[family]:
        if odd(args):
            warnings.warn(message)
