"""Falcon (megatron/model/falcon_model.py:10-41 surface): rotary, parallel attention + MLP,
GeLU, no biases, GQA/MQA via num_attention_heads_kv."""
import warnings

from .. import global_vars
from .enums import PositionEmbeddingType
from .gpt_model import GPTModel


class FalconModel(GPTModel):
    def __init__(self, num_tokentypes: int = 0, parallel_output: bool = True, pre_process: bool = True,
                 post_process: bool = True, model_type=None, args=None):
        args = args or global_vars.get_args()
        assert args.position_embedding_type == PositionEmbeddingType.rotary, \
            f"Falcon uses rotary embedding, not {args.position_embedding_type}"
        assert isinstance(args.num_attention_heads_kv, int), "Falcon needs an explicit num_attention_heads_kv"
        assert not args.use_post_ln, "FalconModel requires pre-normalization, not use_post_ln"
        assert args.glu_activation is None, "FalconModel requires gelu activation (set glu_activation=None)"
        assert not args.use_bias, "Falcon does not use bias"
        assert args.parallel_attn, "Falcon uses parallel_attn"
        if not args.parallel_layernorm:
            warnings.warn("Falcon uses parallel_layernorm, or are you running falcon-7b?")
        if not args.use_flash_attn:
            warnings.warn("Falcon trains fastest with the fused attention kernel (--use_flash_attn)")
        if args.bias_gelu_fusion:
            warnings.warn("Falcon has no MLP bias, so bias_gelu_fusion has no effect")
        if args.bias_dropout_fusion:
            warnings.warn("Falcon has no biases, so bias_dropout_fusion has no effect")
        if args.hidden_dropout > 0.0 and not args.lima_dropout:
            warnings.warn("Falcon is not intended to use dropout")
        super().__init__(num_tokentypes=num_tokentypes, parallel_output=parallel_output,
                         pre_process=pre_process, post_process=post_process, model_type=model_type,
                         args=args)
