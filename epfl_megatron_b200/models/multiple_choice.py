"""Multiple-choice head on the BERT encoder (RACE) (megatron/model/multiple_choice.py:18-119)."""
from __future__ import annotations

import torch

from .. import global_vars
from ._heads import HeadedModel
from .bert_model import bert_extended_attention_mask, bert_position_ids
from .enums import AttnMaskType
from .language_model import get_language_model
from .utils import get_linear_layer, init_method_normal, scaled_init_method_normal


class MultipleChoice(HeadedModel):
    _head_keys = {"multichoice_head": "multichoice_head"}
    _heads_may_be_missing = True

    def __init__(self, num_tokentypes=2, pre_process=True, post_process=True):
        super().__init__(share_word_embeddings=False)
        args = global_vars.get_args()
        self.pre_process, self.post_process = pre_process, post_process
        init_method = init_method_normal(args.init_method_std)
        self.language_model, self._language_model_key = get_language_model(
            num_tokentypes=num_tokentypes, add_pooler=True, encoder_attn_mask_type=AttnMaskType.padding,
            init_method=init_method,
            scaled_init_method=scaled_init_method_normal(args.init_method_std, args.num_layers),
            pre_process=pre_process, post_process=post_process, args=args)
        if post_process:
            self.multichoice_dropout = torch.nn.Dropout(args.hidden_dropout)
            self.multichoice_head = get_linear_layer(args.hidden_size, 1, init_method,
                                                     args.perform_initialization)

    def forward(self, model_input, attention_mask, tokentype_ids=None):
        # inputs are [batch, choices, seq]: fold the choices into the batch, score, unfold
        num_choices = attention_mask.shape[1]
        attention_mask = attention_mask.view(-1, attention_mask.size(-1))
        ext = bert_extended_attention_mask(attention_mask)
        if self.pre_process:
            assert model_input.dim() == 3
            model_input = model_input.view(-1, model_input.size(-1))
            if tokentype_ids is not None:
                tokentype_ids = tokentype_ids.view(-1, tokentype_ids.size(-1))
        out = self.language_model(model_input, bert_position_ids(model_input) if self.pre_process else None,
                                  ext, tokentype_ids=tokentype_ids)
        if not self.post_process:
            return out
        _, pooled = out
        logits = self.multichoice_head(self.multichoice_dropout(pooled))
        return logits.view(-1, num_choices)
