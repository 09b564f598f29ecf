"""T5: encoder-decoder with a tied-embedding LM head (megatron/model/t5_model.py:20-199 surface)."""
from __future__ import annotations

import torch

from .. import global_vars
from ..parallel import layers as tp_layers
from ..parallel.cross_entropy import vocab_parallel_cross_entropy
from ._heads import HeadedModel
from .enums import AttnMaskType
from .language_model import get_language_model, parallel_lm_logits
from .module import MegatronModule
from .utils import init_method_normal, scaled_init_method_normal


def t5_extended_attention_mask(attention_mask_list):
    """[b, sq, sk] masks -> [b, 1, sq, sk] (None passes through)."""
    return [m.unsqueeze(1) if m is not None else None for m in attention_mask_list]


def t5_position_ids(token_ids):
    s = token_ids.size(1)
    return torch.arange(s, dtype=torch.long, device=token_ids.device).unsqueeze(0).expand_as(token_ids)


class T5LMHead(MegatronModule):
    """Logits against the word embeddings plus a vocab-parallel bias."""

    def __init__(self, mpu_vocab_size, parallel_output):
        super().__init__()
        self.bias = torch.nn.Parameter(torch.zeros(mpu_vocab_size))
        tp_layers.set_tensor_model_parallel_attributes(self.bias, True, 0, 1)
        self.parallel_output = parallel_output

    def forward(self, hidden_states, word_embeddings_weight):
        return parallel_lm_logits(hidden_states, word_embeddings_weight, self.parallel_output, bias=self.bias)


class T5Model(HeadedModel):
    _head_keys = {"lm_head": "lm_head"}
    _word_embeddings_for_head_key = "word_embeddings_for_head"

    def __init__(self, num_tokentypes=0, parallel_output=True, pre_process=True, post_process=True,
                 add_encoder=True, add_decoder=True, model_type=None):
        super().__init__()
        args = global_vars.get_args()
        self.fp16_lm_cross_entropy = args.fp16_lm_cross_entropy
        self.parallel_output = parallel_output
        self.pre_process, self.post_process = pre_process, post_process
        self.add_encoder, self.add_decoder = add_encoder, add_decoder
        init_method = init_method_normal(args.init_method_std)
        scaled = scaled_init_method_normal(args.init_method_std, args.num_layers)
        self.language_model, self._language_model_key = get_language_model(
            num_tokentypes=num_tokentypes, add_pooler=False, add_encoder=add_encoder, add_decoder=add_decoder,
            encoder_attn_mask_type=AttnMaskType.padding, init_method=init_method, scaled_init_method=scaled,
            pre_process=pre_process, post_process=post_process, args=args, model_type=model_type)
        self.initialize_word_embeddings(init_method_normal, args)
        self.lm_head = None
        if self.post_process and self.add_decoder:
            self.lm_head = T5LMHead(self.word_embeddings_weight().size(0), parallel_output)

    def _ties_embeddings_on_last_stage(self):
        return self.add_decoder

    def forward(self, encoder_input_ids, decoder_input_ids, encoder_attn_mask, decoder_attn_mask,
                encoder_decoder_attn_mask, tokentype_ids=None, lm_labels=None, enc_hidden_states=None):
        enc_mask, dec_mask, cross_mask = t5_extended_attention_mask(
            [encoder_attn_mask, decoder_attn_mask, encoder_decoder_attn_mask])
        lm_output = self.language_model(encoder_input_ids, t5_position_ids(encoder_input_ids), enc_mask,
                                        decoder_input_ids, t5_position_ids(decoder_input_ids), dec_mask,
                                        cross_mask, tokentype_ids=tokentype_ids,
                                        enc_hidden_states=enc_hidden_states)
        if self.post_process and self.add_decoder:
            decoder_output, _ = lm_output
            logits = self.lm_head(decoder_output, self.word_embeddings_weight())
            if lm_labels is None:
                return logits.transpose(0, 1).contiguous()
            labels = lm_labels.transpose(0, 1).contiguous()
            if self.fp16_lm_cross_entropy:
                assert logits.dtype == torch.half
                loss = vocab_parallel_cross_entropy(logits, labels)
            else:
                loss = vocab_parallel_cross_entropy(logits.float(), labels)
            return loss.transpose(0, 1).contiguous()
        if self.add_decoder and not self.add_encoder:
            decoder_output, _ = lm_output
            return decoder_output
        return lm_output
