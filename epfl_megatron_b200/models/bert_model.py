"""BERT: bidirectional encoder + masked-LM head (+ optional sentence-order binary head)
(megatron/model/bert_model.py:21-243 surface)."""
from __future__ import annotations

import torch

from .. import global_vars
from ..parallel.cross_entropy import vocab_parallel_cross_entropy
from ._heads import HeadedModel, MaskedLMHead
from .enums import AttnMaskType
from .language_model import get_language_model
from .utils import get_linear_layer, init_method_normal, scaled_init_method_normal


def bert_extended_attention_mask(attention_mask):
    """[b, s] 1/0 padding mask -> [b, 1, s, s] bool, True = masked out."""
    m = attention_mask.unsqueeze(1) * attention_mask.unsqueeze(2)
    return (m.unsqueeze(1) < 0.5)


def bert_position_ids(token_ids):
    s = token_ids.size(1)
    return torch.arange(s, dtype=torch.long, device=token_ids.device).unsqueeze(0).expand_as(token_ids)


BertLMHead = MaskedLMHead


def post_language_model_processing(lm_output, pooled_output, lm_head, binary_head, lm_labels, logit_weights,
                                   fp16_lm_cross_entropy):
    lm_logits = lm_head(lm_output, logit_weights)
    binary_logits = binary_head(pooled_output) if binary_head is not None else None
    if lm_labels is None:
        return lm_logits.transpose(0, 1).contiguous(), binary_logits
    lm_labels = lm_labels.transpose(0, 1).contiguous()
    if fp16_lm_cross_entropy:
        assert lm_logits.dtype == torch.half
        loss = vocab_parallel_cross_entropy(lm_logits, lm_labels)
    else:
        loss = vocab_parallel_cross_entropy(lm_logits.float(), lm_labels)
    return loss.transpose(0, 1).contiguous(), binary_logits


class BertModel(HeadedModel):
    _head_keys = {"lm_head": "lm_head", "binary_head": "binary_head"}
    _word_embeddings_for_head_key = "word_embeddings_for_head"

    def __init__(self, num_tokentypes=2, add_binary_head=True, parallel_output=True, pre_process=True,
                 post_process=True, model_type=None):
        super().__init__()
        args = global_vars.get_args()
        self.fp16_lm_cross_entropy = args.fp16_lm_cross_entropy
        self.add_binary_head = add_binary_head
        self.parallel_output = parallel_output
        self.pre_process, self.post_process = pre_process, post_process
        init_method = init_method_normal(args.init_method_std)
        scaled = scaled_init_method_normal(args.init_method_std, args.num_layers)
        self.language_model, self._language_model_key = get_language_model(
            num_tokentypes=num_tokentypes, add_pooler=add_binary_head,
            encoder_attn_mask_type=AttnMaskType.padding, init_method=init_method, scaled_init_method=scaled,
            pre_process=pre_process, post_process=post_process, args=args, model_type=model_type)
        self.initialize_word_embeddings(init_method_normal, args)
        self.lm_head = self.binary_head = None
        if post_process:
            self.lm_head = MaskedLMHead(self.word_embeddings_weight().size(0), args.hidden_size, init_method,
                                        args.layernorm_epsilon, parallel_output)
            if add_binary_head:
                self.binary_head = get_linear_layer(args.hidden_size, 2, init_method,
                                                    args.perform_initialization)

    def _ties_embeddings_on_last_stage(self):
        return True

    def forward(self, bert_model_input, attention_mask, tokentype_ids=None, lm_labels=None):
        ext_mask = bert_extended_attention_mask(attention_mask)
        lm_output = self.language_model(bert_model_input, bert_position_ids(bert_model_input), ext_mask,
                                        tokentype_ids=tokentype_ids)
        if self.post_process and self.add_binary_head:
            lm_output, pooled = lm_output
        else:
            pooled = None
        if not self.post_process:
            return lm_output
        return post_language_model_processing(lm_output, pooled, self.lm_head, self.binary_head, lm_labels,
                                              self.word_embeddings_weight(), self.fp16_lm_cross_entropy)
