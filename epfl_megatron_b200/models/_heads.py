"""Shared pieces of the encoder-model heads: state-dict plumbing for `language_model` plus a set
of named sub-modules, and the masked LM head."""
from __future__ import annotations

import torch

from .. import global_vars
from ..parallel import layers as tp_layers
from .fused_layer_norm import MixedFusedLayerNorm as LayerNorm
from .language_model import parallel_lm_logits
from .module import MegatronModule
from .utils import erf_gelu, get_linear_layer, openai_gelu


class HeadedModel(MegatronModule):
    """A language model plus named heads. Sub-classes list their heads in ``_head_keys``
    (attribute name -> checkpoint key); the checkpoint dict is
    {language_model: ..., <key>: head.state_dict(), [word_embeddings_for_head: ...]}."""
    _language_model_key = "language_model"
    _head_keys: dict = {}
    # task heads put on top of a pretrained model start from random weights when the checkpoint has
    # none (megatron/model/classification.py:97-108); pretraining heads must be in the checkpoint
    _heads_may_be_missing = False

    def set_input_tensor(self, input_tensor):
        self.language_model.set_input_tensor(input_tensor)

    def _ties_embeddings_on_last_stage(self):
        return False

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        sd = {self._language_model_key:
              self.language_model.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)}
        if self.post_process:
            for attr, key in self._head_keys.items():
                mod = getattr(self, attr, None)
                if mod is not None:
                    sd[key] = mod.state_dict(prefix=prefix, keep_vars=keep_vars)
        if self._ties_embeddings_on_last_stage() and self.post_process and not self.pre_process:
            sd[self._word_embeddings_for_head_key] = self.word_embeddings.state_dict(
                prefix=prefix, keep_vars=keep_vars)
        return sd

    def load_state_dict(self, state_dict, strict=True):
        self.language_model.load_state_dict(state_dict[self._language_model_key], strict=strict)
        if self.post_process:
            for attr, key in self._head_keys.items():
                mod = getattr(self, attr, None)
                if mod is not None and key in state_dict:
                    mod.load_state_dict(state_dict[key], strict=strict)
                elif mod is not None and self._heads_may_be_missing:
                    print(f"***WARNING*** could not find {key} in the checkpoint, initializing to random",
                          flush=True)
                elif mod is not None:
                    raise KeyError(f"{key} is missing from the checkpoint")
        if self._ties_embeddings_on_last_stage() and self.post_process and not self.pre_process:
            self.word_embeddings.load_state_dict(state_dict[self._word_embeddings_for_head_key], strict=strict)


class MaskedLMHead(MegatronModule):
    """dense -> gelu -> layernorm -> logits against the (tied) word embeddings + vocab bias
    (megatron/model/bert_model.py:47-91)."""

    def __init__(self, mpu_vocab_size, hidden_size, init_method, layernorm_epsilon, parallel_output):
        super().__init__()
        args = global_vars.get_args()
        self.bias = torch.nn.Parameter(torch.zeros(mpu_vocab_size))
        tp_layers.set_tensor_model_parallel_attributes(self.bias, True, 0, 1)
        self.parallel_output = parallel_output
        self.dense = get_linear_layer(hidden_size, hidden_size, init_method,
                                      perform_initialization=args.perform_initialization)
        for p in (self.dense.weight, self.dense.bias):
            setattr(p, "sequence_parallel", args.sequence_parallel)
        self.layernorm = LayerNorm(hidden_size, eps=layernorm_epsilon, sequence_parallel=args.sequence_parallel)
        self.gelu = torch.nn.functional.gelu
        if getattr(args, "openai_gelu", False):
            self.gelu = openai_gelu
        elif args.onnx_safe:
            self.gelu = erf_gelu

    def forward(self, hidden_states, word_embeddings_weight):
        h = self.layernorm(self.gelu(self.dense(hidden_states)))
        return parallel_lm_logits(h, word_embeddings_weight, self.parallel_output, bias=self.bias)
