"""Sequence classification on top of the BERT encoder (GLUE) (megatron/model/classification.py:17-106)."""
from __future__ import annotations

import torch

from .. import global_vars
from ..utils import print_rank_last
from ._heads import HeadedModel
from .bert_model import bert_extended_attention_mask, bert_position_ids
from .enums import AttnMaskType
from .language_model import get_language_model
from .utils import get_linear_layer, init_method_normal, scaled_init_method_normal


class Classification(HeadedModel):
    _head_keys = {"classification_head": "classification_head"}
    _heads_may_be_missing = True

    def __init__(self, num_classes, num_tokentypes=2, pre_process=True, post_process=True):
        super().__init__(share_word_embeddings=False)
        args = global_vars.get_args()
        self.num_classes = num_classes
        self.pre_process, self.post_process = pre_process, post_process
        init_method = init_method_normal(args.init_method_std)
        self.language_model, self._language_model_key = get_language_model(
            num_tokentypes=num_tokentypes, add_pooler=True, encoder_attn_mask_type=AttnMaskType.padding,
            init_method=init_method,
            scaled_init_method=scaled_init_method_normal(args.init_method_std, args.num_layers),
            pre_process=pre_process, post_process=post_process, args=args)
        if post_process:
            self.classification_dropout = torch.nn.Dropout(args.hidden_dropout)
            self.classification_head = get_linear_layer(args.hidden_size, num_classes, init_method,
                                                        args.perform_initialization)

    def forward(self, model_input, attention_mask, tokentype_ids=None):
        ext = bert_extended_attention_mask(attention_mask)
        out = self.language_model(model_input, bert_position_ids(model_input), ext, tokentype_ids=tokentype_ids)
        if not self.post_process:
            return out
        _, pooled = out
        logits = self.classification_head(self.classification_dropout(pooled))
        return logits.view(-1, self.num_classes)

    def load_state_dict(self, state_dict, strict=True):
        if self.post_process and "classification_head" not in state_dict:
            print_rank_last("***WARNING*** could not find classification_head in the checkpoint, "
                            "initializing to random")
        super().load_state_dict(state_dict, strict=strict)
