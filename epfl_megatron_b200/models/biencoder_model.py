"""Dual-encoder retriever (ICT / ORQA): query tower + context tower of BERT encoders, optionally
sharing weights, each with an optional projection (megatron/model/biencoder_model.py:19-349)."""
from __future__ import annotations

import os
import sys

import torch

from .. import global_vars
from ..checkpointing import fix_query_key_value_ordering, get_checkpoint_names, get_checkpoint_tracker_filename
from ..parallel import state as mpu
from ..utils import print_rank_0
from .bert_model import bert_extended_attention_mask, bert_position_ids
from .enums import AttnMaskType, ModelType
from .language_model import get_language_model
from .module import MegatronModule
from .utils import get_linear_layer, init_method_normal, scaled_init_method_normal


def get_model_provider(only_query_model=False, only_context_model=False, biencoder_shared_query_context_model=False):
    def model_provider(pre_process=True, post_process=True):
        print_rank_0("building Bienoder model ...")
        return biencoder_model_provider(
            only_query_model=only_query_model, only_context_model=only_context_model,
            biencoder_shared_query_context_model=biencoder_shared_query_context_model,
            pre_process=pre_process, post_process=post_process)
    return model_provider


def biencoder_model_provider(only_query_model=False, only_context_model=False,
                             biencoder_shared_query_context_model=False, pre_process=True, post_process=True):
    assert mpu.get_tensor_model_parallel_world_size() == 1 and \
        mpu.get_pipeline_model_parallel_world_size() == 1, "model parallel size > 1 not supported for ICT"
    print_rank_0("building BiEncoderModel...")
    return BiEncoderModel(num_tokentypes=2, parallel_output=False, only_query_model=only_query_model,
                          only_context_model=only_context_model,
                          biencoder_shared_query_context_model=biencoder_shared_query_context_model,
                          pre_process=pre_process, post_process=post_process)


class BiEncoderModel(MegatronModule):
    def __init__(self, num_tokentypes=1, parallel_output=True, only_query_model=False,
                 only_context_model=False, biencoder_shared_query_context_model=False,
                 pre_process=True, post_process=True):
        super().__init__()
        args = global_vars.get_args()
        kw = dict(num_tokentypes=num_tokentypes, parallel_output=parallel_output, pre_process=pre_process,
                  post_process=post_process)
        self.biencoder_shared_query_context_model = biencoder_shared_query_context_model
        assert not (only_context_model and only_query_model)
        self.use_context_model = not only_query_model
        self.use_query_model = not only_context_model
        self.biencoder_projection_dim = args.biencoder_projection_dim
        if self.biencoder_shared_query_context_model:
            self.model = PretrainedBertModel(**kw)
            self._model_key = "shared_model"
            self.query_model, self.context_model = self.model, self.model
        else:
            if self.use_query_model:
                self.query_model = PretrainedBertModel(**kw)
                self._query_key = "query_model"
            if self.use_context_model:
                self.context_model = PretrainedBertModel(**kw)
                self._context_key = "context_model"

    def set_input_tensor(self, input_tensor):
        # one stage only: nothing is ever received from a previous stage
        return

    def forward(self, query_tokens, query_attention_mask, query_types, context_tokens,
                context_attention_mask, context_types):
        """-> (query_logits, context_logits)"""
        if self.use_query_model:
            q = self.embed_text(self.query_model, query_tokens, query_attention_mask, query_types)
        else:
            raise ValueError("Cannot embed query without the query model.")
        if self.use_context_model:
            c = self.embed_text(self.context_model, context_tokens, context_attention_mask, context_types)
        else:
            raise ValueError("Cannot embed block without the block model.")
        return q, c

    @staticmethod
    def embed_text(model, tokens, attention_mask, token_types):
        return model(tokens, attention_mask, token_types)

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        sd = {}
        if self.biencoder_shared_query_context_model:
            sd[self._model_key] = self.model.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
        else:
            if self.use_query_model:
                sd[self._query_key] = self.query_model.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
            if self.use_context_model:
                sd[self._context_key] = self.context_model.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
        return sd

    def load_state_dict(self, state_dict, strict=True):
        if self.biencoder_shared_query_context_model:
            print_rank_0("Loading shared query-context model")
            self.model.load_state_dict(state_dict[self._model_key], strict=strict)
        else:
            if self.use_query_model:
                print_rank_0("Loading query model")
                self.query_model.load_state_dict(state_dict[self._query_key], strict=strict)
            if self.use_context_model:
                print_rank_0("Loading context model")
                self.context_model.load_state_dict(state_dict[self._context_key], strict=strict)

    def init_state_dict_from_bert(self):
        """Warm-start both towers from a pretrained BERT checkpoint (--bert_load)."""
        args = global_vars.get_args()
        if args.bert_load is None:
            print_rank_0("bert_load is None, not initializing from a pretrained BERT")
            return
        tracker = get_checkpoint_tracker_filename(args.bert_load)
        if not os.path.isfile(tracker):
            raise FileNotFoundError("Could not find BERT checkpoint")
        with open(tracker) as f:
            iteration = int(f.read().strip())
            assert iteration > 0
        name, _ = get_checkpoint_names(args.bert_load, iteration, False)
        if mpu.get_data_parallel_rank() == 0:
            print(f"global rank {torch.distributed.get_rank()} is loading BERT checkpoint {name}")
        from ..checkpointing import _torch_load
        state_dict = _torch_load(name)
        model_dict = state_dict["model"]["language_model"]
        if self.biencoder_shared_query_context_model:
            self.model.language_model.load_state_dict(model_dict)
            fix_query_key_value_ordering(self.model, state_dict.get("checkpoint_version", 0))
        else:
            if self.use_query_model:
                self.query_model.language_model.load_state_dict(model_dict)
                fix_query_key_value_ordering(self.query_model, state_dict.get("checkpoint_version", 0))
            if self.use_context_model:
                self.context_model.language_model.load_state_dict(model_dict)
                fix_query_key_value_ordering(self.context_model, state_dict.get("checkpoint_version", 0))


class PretrainedBertModel(MegatronModule):
    """BERT encoder without LM head: returns the [CLS] vector, optionally projected."""

    def __init__(self, num_tokentypes=2, parallel_output=True, pre_process=True, post_process=True):
        super().__init__()
        args = global_vars.get_args()
        tokenizer = global_vars.get_tokenizer()
        self.pad_id = tokenizer.pad
        self.biencoder_projection_dim = args.biencoder_projection_dim
        self.parallel_output = parallel_output
        self.pre_process, self.post_process = pre_process, post_process
        init_method = init_method_normal(args.init_method_std)
        self.language_model, self._language_model_key = get_language_model(
            num_tokentypes=num_tokentypes, add_pooler=False, encoder_attn_mask_type=AttnMaskType.padding,
            init_method=init_method,
            scaled_init_method=scaled_init_method_normal(args.init_method_std, args.num_layers),
            pre_process=pre_process, post_process=post_process, args=args)
        if args.biencoder_projection_dim > 0:
            self.projection_enc = get_linear_layer(args.hidden_size, args.biencoder_projection_dim,
                                                   init_method, args.perform_initialization)
            self._projection_enc_key = "projection_enc"

    def forward(self, input_ids, attention_mask, tokentype_ids=None):
        ext = attention_mask.unsqueeze(1)       # already [b, s, s] bool, True = masked out
        lm_output = self.language_model(input_ids, bert_position_ids(input_ids), ext, tokentype_ids=tokentype_ids)
        pooled = lm_output[0, :, :]             # [s, b, h] -> the [CLS] row
        pooled = pooled.clone()
        if self.biencoder_projection_dim:
            pooled = self.projection_enc(pooled)
        return pooled

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        sd = {self._language_model_key:
              self.language_model.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)}
        if self.biencoder_projection_dim > 0:
            sd[self._projection_enc_key] = self.projection_enc.state_dict(prefix=prefix, keep_vars=keep_vars)
        return sd

    def load_state_dict(self, state_dict, strict=True):
        print_rank_0("loading pretrained weights")
        self.language_model.load_state_dict(state_dict[self._language_model_key], strict=strict)
        if self.biencoder_projection_dim > 0:
            print_rank_0("loading projection head weights")
            self.projection_enc.load_state_dict(state_dict[self._projection_enc_key], strict=strict)
