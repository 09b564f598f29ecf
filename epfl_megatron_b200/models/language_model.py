"""Embedding + transformer (+ optional decoder / pooler / untied lm_head)
(megatron/model/language_model.py:24-638 surface and state-dict keys)."""
from __future__ import annotations

import torch

from .. import global_vars
from ..parallel import layers as tp_layers
from ..parallel import random as tp_random
from ..parallel import state as mpu
from ..parallel.layers import linear_with_grad_accumulation_and_async_allreduce
from ..parallel.mappings import (copy_to_tensor_model_parallel_region,
                                 gather_from_tensor_model_parallel_region,
                                 reduce_scatter_to_sequence_parallel_region,
                                 reduce_from_tensor_model_parallel_region,
                                 scatter_to_sequence_parallel_region)
from .enums import AttnMaskType, LayerType, PositionEmbeddingType
from .module import MegatronModule
from .transformer import ParallelTransformer
from .utils import get_linear_layer, init_method_normal, scaled_init_method_normal


def parallel_lm_logits(input_, word_embeddings_weight, parallel_output, bias=None, args=None):
    """LM head through the same autograd linear as every other TP GEMM; the output is the
    vocab-parallel logits [s, b, V/tp] (gathered when parallel_output is False)."""
    args = args or global_vars.get_args()
    async_ar = (args.async_tensor_model_parallel_allreduce
                and mpu.get_tensor_model_parallel_world_size() > 1 and not args.sequence_parallel)
    if async_ar or args.sequence_parallel or mpu.get_tensor_model_parallel_world_size() == 1:
        x = input_
    else:
        x = copy_to_tensor_model_parallel_region(input_)
    logits = linear_with_grad_accumulation_and_async_allreduce(
        x, word_embeddings_weight, bias, args.gradient_accumulation_fusion, async_ar,
        args.sequence_parallel)
    if parallel_output:
        return logits
    return gather_from_tensor_model_parallel_region(logits)


def get_language_model(num_tokentypes, add_pooler, encoder_attn_mask_type, init_method=None,
                       scaled_init_method=None, add_encoder=True, add_decoder=False,
                       decoder_attn_mask_type=AttnMaskType.causal, pre_process=True, post_process=True,
                       args=None, model_type=None):
    args = args or global_vars.get_args()
    if init_method is None:
        init_method = init_method_normal(args.init_method_std)
    if scaled_init_method is None:
        scaled_init_method = scaled_init_method_normal(args.init_method_std, args.num_layers)
    lm = TransformerLanguageModel(init_method, scaled_init_method, encoder_attn_mask_type,
                                  num_tokentypes=num_tokentypes, add_encoder=add_encoder,
                                  add_decoder=add_decoder, decoder_attn_mask_type=decoder_attn_mask_type,
                                  add_pooler=add_pooler, pre_process=pre_process, post_process=post_process,
                                  args=args, model_type=model_type)
    return lm, "language_model"


class Pooler(MegatronModule):
    """tanh(W * hidden[sequence_index]) (BERT [CLS] pooling)."""

    def __init__(self, hidden_size, init_method, args):
        super().__init__()
        self.dense = get_linear_layer(hidden_size, hidden_size, init_method, args.perform_initialization)
        self.sequence_parallel = args.sequence_parallel

    def forward(self, hidden_states, sequence_index=0):
        if self.sequence_parallel:
            from ..parallel.mappings import gather_from_sequence_parallel_region
            hidden_states = gather_from_sequence_parallel_region(hidden_states, tensor_parallel_output_grad=False)
        return torch.tanh(self.dense(hidden_states[sequence_index, :, :]))


class Embedding(MegatronModule):
    """Word (+ learned position, + token type) embeddings -> [s, b, h] ([s/tp, b, h] with SP)."""

    def __init__(self, hidden_size, vocab_size, max_sequence_length, embedding_dropout_prob, init_method,
                 num_tokentypes=0, args=None):
        super().__init__()
        args = args or global_vars.get_args()
        self.hidden_size, self.init_method, self.num_tokentypes = hidden_size, init_method, num_tokentypes
        self.word_embeddings = tp_layers.VocabParallelEmbedding(
            vocab_size, hidden_size, init_method=init_method, params_dtype=args.params_dtype,
            use_cpu_initialization=args.use_cpu_initialization,
            perform_initialization=args.perform_initialization)
        self._word_embeddings_key = "word_embeddings"
        self.position_embedding_type = args.position_embedding_type
        if self.position_embedding_type == PositionEmbeddingType.absolute:
            assert max_sequence_length is not None
            self.position_embeddings = torch.nn.Embedding(max_sequence_length, hidden_size)
            if args.perform_initialization:
                init_method(self.position_embeddings.weight)
        else:
            self.position_embeddings = None
        self._position_embeddings_key = "position_embeddings"
        self._tokentype_embeddings_key = "tokentype_embeddings"
        if num_tokentypes > 0:
            self.tokentype_embeddings = torch.nn.Embedding(num_tokentypes, hidden_size)
            if args.perform_initialization:
                init_method(self.tokentype_embeddings.weight)
        else:
            self.tokentype_embeddings = None
        self.fp32_residual_connection = args.fp32_residual_connection
        self.sequence_parallel = args.sequence_parallel
        self.embedding_dropout = torch.nn.Dropout(embedding_dropout_prob)
        self.perform_initialization = args.perform_initialization

    def zero_parameters(self):
        self.word_embeddings.weight.data.fill_(0)
        self.word_embeddings.weight.shared = True
        if self.position_embeddings is not None:
            self.position_embeddings.weight.data.fill_(0)
            self.position_embeddings.weight.shared = True
        if self.num_tokentypes > 0:
            self.tokentype_embeddings.weight.data.fill_(0)
            self.tokentype_embeddings.weight.shared = True

    def add_tokentype_embeddings(self, num_tokentypes):
        if self.tokentype_embeddings is not None:
            raise Exception("tokentype embeddings is already initialized")
        if torch.distributed.is_initialized() and torch.distributed.get_rank() == 0:
            print(f"adding embedding for {num_tokentypes} tokentypes", flush=True)
        self.num_tokentypes = num_tokentypes
        self.tokentype_embeddings = torch.nn.Embedding(num_tokentypes, self.hidden_size)
        self.init_method(self.tokentype_embeddings.weight)

    def forward(self, input_ids, position_ids, tokentype_ids=None):
        tp = mpu.get_tensor_model_parallel_world_size()
        extra = (self.position_embeddings is not None) or (tokentype_ids is not None)
        if not extra and torch.is_grad_enabled() and input_ids.is_cuda and \
                tp_layers.fused_vocab_embedding_ok(input_ids, self.word_embeddings.weight, self.sequence_parallel):
            # rows copied straight from the owning rank's table shard over peer memory: no reduction
            emb = tp_layers._FusedVocabEmbedding.apply(input_ids, self.word_embeddings.weight,
                                                       self.sequence_parallel)
            if self.fp32_residual_connection:
                emb = emb.float()
            if self.sequence_parallel and tp > 1:
                with tp_random.get_cuda_rng_tracker().fork():
                    return self.embedding_dropout(emb)
            return self.embedding_dropout(emb)
        if self.sequence_parallel and tp > 1 and not extra:
            # partial lookup -> [s, b, h] -> reduce-scatter along s: replaces the reference's
            # all-reduce followed by discarding (tp-1)/tp of the result
            part = self.word_embeddings.local_lookup(input_ids).transpose(0, 1).contiguous()
            if self.fp32_residual_connection:
                part = part.float()
            emb = reduce_scatter_to_sequence_parallel_region(part)
            with tp_random.get_cuda_rng_tracker().fork():
                return self.embedding_dropout(emb)
        emb = self.word_embeddings(input_ids)
        if self.position_embeddings is not None:
            emb = emb + self.position_embeddings(position_ids)
        if tokentype_ids is not None:
            assert self.tokentype_embeddings is not None
            emb = emb + self.tokentype_embeddings(tokentype_ids)
        else:
            assert self.tokentype_embeddings is None
        emb = emb.transpose(0, 1).contiguous()
        if self.fp32_residual_connection:
            emb = emb.float()
        if self.sequence_parallel:
            emb = scatter_to_sequence_parallel_region(emb)
            with tp_random.get_cuda_rng_tracker().fork():
                emb = self.embedding_dropout(emb)
        else:
            emb = self.embedding_dropout(emb)
        return emb

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        sd = {self._word_embeddings_key: self.word_embeddings.state_dict(prefix=prefix, keep_vars=keep_vars)}
        if self.position_embeddings is not None:
            sd[self._position_embeddings_key] = self.position_embeddings.state_dict(prefix=prefix, keep_vars=keep_vars)
        if self.num_tokentypes > 0:
            sd[self._tokentype_embeddings_key] = self.tokentype_embeddings.state_dict(prefix=prefix, keep_vars=keep_vars)
        return sd

    def load_state_dict(self, state_dict, strict=True):
        if self._word_embeddings_key in state_dict:
            w = state_dict[self._word_embeddings_key]
        else:  # legacy flat spelling: 'word_embeddings.weight'
            w = {k.split("word_embeddings.")[1]: v for k, v in state_dict.items() if "word_embeddings" in k}
        self.word_embeddings.load_state_dict(w, strict=strict)
        if self.position_embeddings is not None:
            if self._position_embeddings_key in state_dict:
                p = state_dict[self._position_embeddings_key]
            else:
                p = {k.split("position_embeddings.")[1]: v for k, v in state_dict.items()
                     if "position_embeddings" in k}
            self.position_embeddings.load_state_dict(p, strict=strict)
        if self.num_tokentypes > 0:
            if self._tokentype_embeddings_key in state_dict:
                t = state_dict[self._tokentype_embeddings_key]
            else:
                t = {k.split("tokentype_embeddings.")[1]: v for k, v in state_dict.items()
                     if "tokentype_embeddings" in k}
            if t:
                self.tokentype_embeddings.load_state_dict(t, strict=strict)
            else:
                print("***WARNING*** expected tokentype embeddings in the checkpoint but could not find it",
                      flush=True)


class TransformerLanguageModel(MegatronModule):
    def __init__(self, init_method, output_layer_init_method, encoder_attn_mask_type, num_tokentypes=0,
                 add_encoder=True, add_decoder=False, decoder_attn_mask_type=AttnMaskType.causal,
                 add_pooler=False, pre_process=True, post_process=True, args=None, model_type=None):
        super().__init__()
        args = args or global_vars.get_args()
        self.pre_process, self.post_process = pre_process, post_process
        self.hidden_size = args.hidden_size
        self.num_tokentypes = num_tokentypes
        self.init_method = init_method
        self.add_encoder, self.add_decoder = add_encoder, add_decoder
        self.encoder_attn_mask_type, self.decoder_attn_mask_type = encoder_attn_mask_type, decoder_attn_mask_type
        self.add_pooler = add_pooler
        self.encoder_hidden_state = None
        s, ell, h, V = args.seq_length, args.num_layers, args.hidden_size, args.padded_vocab_size
        # per-sequence forward FLOPs: attention projections + scores + MLP, then the head
        self.flop_estimate = 24 * s * ell * h * h * (1 + s / (6 * h)) + 2 * s * h * V
        if self.pre_process:
            self.embedding = Embedding(self.hidden_size, args.padded_vocab_size, args.max_position_embeddings,
                                       args.hidden_dropout, self.init_method, self.num_tokentypes, args)
            self._embedding_key = "embedding"
        if self.add_encoder:
            self.encoder = ParallelTransformer(self.init_method, output_layer_init_method,
                                               self_attn_mask_type=self.encoder_attn_mask_type,
                                               pre_process=self.pre_process, post_process=self.post_process,
                                               args=args, model_type=model_type)
            self._encoder_key = "encoder"
        else:
            self.encoder = None
        if self.add_decoder:
            self.decoder = ParallelTransformer(self.init_method, output_layer_init_method,
                                               layer_type=LayerType.decoder,
                                               self_attn_mask_type=self.decoder_attn_mask_type,
                                               pre_process=self.pre_process, post_process=self.post_process,
                                               args=args, model_type=model_type)
            self._decoder_key = "decoder"
        else:
            self.decoder = None
        if self.post_process:
            if self.add_pooler:
                self.pooler = Pooler(self.hidden_size, self.init_method, args)
                self._pooler_key = "pooler"
            self.tie_embed_logits = args.tie_embed_logits
            if not self.tie_embed_logits:
                # untied head: a raw [V/tp, h] Parameter saved under the 'lm_head' key
                tp = mpu.get_tensor_model_parallel_world_size()
                rows = args.padded_vocab_size // tp
                dev = "cpu" if (args.use_cpu_initialization or not torch.cuda.is_available()) \
                    else torch.cuda.current_device()
                self.lm_head = torch.nn.Parameter(torch.empty(rows, self.hidden_size, device=dev,
                                                              dtype=args.params_dtype))
                # Xavier init over the FULL [V, h] matrix (reference: language_model.py:448)
                head_init = torch.nn.init.xavier_uniform_ if args.init_method_xavier_uniform \
                    else torch.nn.init.xavier_normal_
                if args.perform_initialization:
                    if dev == "cpu":
                        master = torch.empty(args.padded_vocab_size, self.hidden_size, dtype=torch.float)
                        head_init(master)
                        r = mpu.get_tensor_model_parallel_rank()
                        with torch.no_grad():
                            self.lm_head.copy_(master[r * rows:(r + 1) * rows].to(args.params_dtype))
                        tp_layers.set_tensor_model_parallel_attributes(self.lm_head, True, 0, 1)
                    else:
                        # a shard must be drawn with the fan of the full matrix, not of the shard
                        std = (2.0 / (args.padded_vocab_size + self.hidden_size)) ** 0.5
                        if args.init_method_xavier_uniform:
                            bound = std * 3.0 ** 0.5
                            shard_init = lambda w: torch.nn.init.uniform_(w, -bound, bound)  # noqa: E731
                        else:
                            shard_init = lambda w: torch.nn.init.normal_(w, 0.0, std)  # noqa: E731
                        tp_layers._initialize_affine_weight_gpu(self.lm_head, shard_init, 0, 1)
                else:
                    tp_layers.set_tensor_model_parallel_attributes(self.lm_head, True, 0, 1)
                self._lm_key = "lm_head"
            else:
                self.lm_head = None

    def set_input_tensor(self, input_tensor):
        if not isinstance(input_tensor, list):
            input_tensor = [input_tensor]
        if self.add_encoder and self.add_decoder:
            assert len(input_tensor) == 1
            self.encoder.set_input_tensor(input_tensor[0])
        elif self.add_encoder:
            assert len(input_tensor) == 1
            self.encoder.set_input_tensor(input_tensor[0])
        elif self.add_decoder:
            if len(input_tensor) == 2:
                self.decoder.set_input_tensor(input_tensor[0])
                self.encoder_hidden_state = input_tensor[1]
            elif len(input_tensor) == 1:
                self.decoder.set_input_tensor(None)
                self.encoder_hidden_state = input_tensor[0]
            else:
                raise Exception("input_tensor must have either length 1 or 2")
        else:
            raise Exception("Stage must have at least either encoder or decoder")

    def forward(self, enc_input_ids, enc_position_ids, enc_attn_mask, dec_input_ids=None,
                dec_position_ids=None, dec_attn_mask=None, enc_dec_attn_mask=None, tokentype_ids=None,
                inference_params=None, pooling_sequence_index=0, enc_hidden_states=None,
                output_enc_hidden=False):
        encoder_input = None
        if self.pre_process:
            encoder_input = self.embedding(enc_input_ids, enc_position_ids, tokentype_ids=tokentype_ids)
        if enc_hidden_states is None:
            if self.encoder is not None:
                encoder_output = self.encoder(encoder_input, enc_attn_mask, inference_params=inference_params,
                                              position_ids=enc_position_ids)
            else:
                encoder_output = self.encoder_hidden_state
        else:
            encoder_output = enc_hidden_states.to(encoder_input.dtype)
        pooled = None
        if self.post_process and self.add_pooler:
            pooled = self.pooler(encoder_output, pooling_sequence_index)
        if not self.add_decoder or output_enc_hidden:
            if self.add_pooler and self.post_process:
                return encoder_output, pooled
            return encoder_output
        decoder_input = self.embedding(dec_input_ids, dec_position_ids) if self.pre_process else None
        decoder_output = self.decoder(decoder_input, dec_attn_mask, encoder_output=encoder_output,
                                      enc_dec_attn_mask=enc_dec_attn_mask, inference_params=inference_params)
        if self.add_pooler and self.post_process:
            return decoder_output, encoder_output, pooled
        return decoder_output, encoder_output

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        sd = {}
        if self.pre_process:
            sd[self._embedding_key] = self.embedding.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
        if self.add_encoder:
            sd[self._encoder_key] = self.encoder.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
        if self.post_process:
            if self.add_pooler:
                sd[self._pooler_key] = self.pooler.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
            if not self.tie_embed_logits:
                sd[self._lm_key] = self.lm_head.data
        if self.add_decoder:
            sd[self._decoder_key] = self.decoder.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)
        return sd

    def load_state_dict(self, state_dict, strict=True):
        if self.pre_process:
            if self._embedding_key in state_dict:
                e = state_dict[self._embedding_key]
            else:
                e = {k: v for k, v in state_dict.items() if "_embeddings" in k}
            self.embedding.load_state_dict(e, strict=strict)
        if self.post_process and not self.tie_embed_logits:
            self.lm_head.data.copy_(state_dict[self._lm_key])
        if self.add_encoder:
            if self._encoder_key in state_dict:
                t = state_dict[self._encoder_key]
            elif "transformer" in state_dict:      # legacy alias written by weights2megatron
                t = state_dict["transformer"]
            else:
                t = {k.split("transformer.")[1]: v for k, v in state_dict.items() if "transformer." in k}
            fixed = {}
            for k, v in t.items():
                fixed[k.replace(".attention.", ".self_attention.") if ".attention." in k else k] = v
            self.encoder.load_state_dict(fixed, strict=strict)
        if self.post_process and self.add_pooler:
            assert "pooler" in state_dict, "could not find data for pooler in the checkpoint"
            self.pooler.load_state_dict(state_dict[self._pooler_key], strict=strict)
        if self.add_decoder:
            assert "decoder" in state_dict, "could not find data for decoder in the checkpoint"
            self.decoder.load_state_dict(state_dict[self._decoder_key], strict=strict)
