"""Transformer stack (megatron/model/transformer.py:77-1251 surface and parameter names).

Notation: s sequence, b micro-batch, h hidden, np heads per TP rank, g KV groups per TP
rank, hn head dim. Activations are [s, b, h] ([s/tp, b, h] under sequence parallelism).

What is different inside (B200-first):
* attention consumes the fused QKV projection output [s, b, g, q_per_g+2, hn] directly:
  RoPE is applied in place on it and the tcgen05 attention kernel reads Q/K/V through
  strided TMA maps -- no K/V head expansion, no [s,b]<->[b,s] copies
  (reference: transformer.py:448-455, 515-522);
* residual add + RMSNorm are one kernel; SwiGLU / GeLU / norms / dropout-add are single
  kernels instead of eager op chains;
* the RoPE table lives on the device.
"""
from __future__ import annotations

import math
from contextlib import nullcontext

import torch
import torch.nn.functional as F

from .. import global_vars
from ..ops import functional as Fn
from ..ops import ref as ops_ref
from ..parallel import layers as tp_layers
from ..parallel import random as tp_random
from ..parallel import state as mpu
from ..parallel.mappings import (gather_from_sequence_parallel_region,
                                 scatter_to_sequence_parallel_region)
from ..parallel.utils import divide, make_viewless_tensor
from .enums import AttnMaskType, AttnType, LayerType, ModelType, PositionEmbeddingType
from .fused_layer_norm import MixedFusedLayerNorm as LayerNorm
from .fused_layer_norm import RMSNorm
from .fused_softmax import FusedScaleMaskSoftmax
from .glu_activations import GLU_ACTIVATIONS
from .module import MegatronModule
from .positional_embeddings import precompute_freqs_cis
from .utils import attention_mask_func, erf_gelu


def _args_to_kwargs(args):
    return dict(params_dtype=args.params_dtype, use_cpu_initialization=args.use_cpu_initialization,
                perform_initialization=args.perform_initialization,
                gradient_accumulation_fusion=args.gradient_accumulation_fusion,
                sequence_parallel_enabled=args.sequence_parallel)


class DropPath(MegatronModule):
    def __init__(self, drop_prob=0.0):
        super().__init__()
        self.drop_prob = drop_prob

    def forward(self, hidden_state):
        if self.drop_prob == 0.0 or not self.training:
            return hidden_state
        keep = 1 - self.drop_prob
        shape = (1,) + (hidden_state.shape[1],) + (1,) * (hidden_state.ndim - 2)
        mask = (keep + torch.rand(shape, dtype=hidden_state.dtype, device=hidden_state.device)).floor_()
        return hidden_state.div(keep) * mask


class ParallelMLP(MegatronModule):
    """h -> ffn (2*ffn for GLU) -> activation -> h."""

    def __init__(self, init_method, output_layer_init_method, args, world_size=None):
        super().__init__()
        self.add_bias = args.use_bias
        self.glu_kind = args.glu_activation
        ffn_in = args.ffn_hidden_size * 2 if args.glu_activation else args.ffn_hidden_size
        kw = _args_to_kwargs(args)
        self.dense_h_to_4h = tp_layers.ColumnParallelLinear(
            args.hidden_size, ffn_in, bias=self.add_bias, gather_output=False, init_method=init_method,
            skip_bias_add=True, async_tensor_model_parallel_allreduce=args.async_tensor_model_parallel_allreduce,
            world_size=world_size,
            # each rank holds [up_shard; gate_shard]: stride 2 makes host-side init independent of tp
            stride=2 if args.glu_activation else 1, **kw)
        self.bias_gelu_fusion = args.bias_gelu_fusion
        self.onnx_safe = args.onnx_safe
        self.dense_4h_to_h = tp_layers.RowParallelLinear(
            args.ffn_hidden_size, args.hidden_size, bias=self.add_bias, input_is_parallel=True,
            init_method=output_layer_init_method, skip_bias_add=True, world_size=world_size, **kw)

    def _activation(self, x, bias):
        if self.glu_kind:
            if bias is not None:
                x = x + bias
            return Fn.glu(x, self.glu_kind)
        if self.onnx_safe:
            return erf_gelu(x if bias is None else x + bias)
        # tanh flavour when fused with a bias (reference jit fusion), exact erf otherwise
        return Fn.bias_gelu(x, bias, tanh_approx=self.bias_gelu_fusion and bias is not None)

    def forward(self, hidden_states):
        inter, bias = self.dense_h_to_4h(hidden_states)
        inter = self._activation(inter, bias)
        return self.dense_4h_to_h(inter)


class CoreAttention(MegatronModule):
    """Unfused attention: QK^T -> scale/mask/softmax -> dropout -> PV (BERT/T5 padding masks,
    inference with a KV cache, head dims the fused kernel does not cover)."""

    def __init__(self, layer_number, args, attn_mask_type=AttnMaskType.padding, world_size=None):
        super().__init__()
        self.fp16, self.bf16 = args.fp16, args.bf16
        self.apply_query_key_layer_scaling = args.apply_query_key_layer_scaling
        self.attention_softmax_in_fp32 = args.attention_softmax_in_fp32
        if self.apply_query_key_layer_scaling:
            self.attention_softmax_in_fp32 = True
        self.layer_number = max(1, layer_number)
        self.attn_mask_type = attn_mask_type
        self.sequence_parallel = args.sequence_parallel
        proj = args.kv_channels * args.num_attention_heads
        ws = world_size if world_size is not None else mpu.get_tensor_model_parallel_world_size()
        self.hidden_size_per_partition = divide(proj, ws)
        self.hidden_size_per_attention_head = divide(proj, args.num_attention_heads)
        self.num_attention_heads_per_partition = divide(args.num_attention_heads, ws)
        coeff = None
        self.norm_factor = math.sqrt(self.hidden_size_per_attention_head)
        if self.apply_query_key_layer_scaling:
            coeff = self.layer_number
            self.norm_factor *= coeff
        self.scale_mask_softmax = FusedScaleMaskSoftmax(
            self.fp16, self.bf16, self.attn_mask_type, args.masked_softmax_fusion, attention_mask_func,
            self.attention_softmax_in_fp32, coeff)
        self.attention_dropout = torch.nn.Dropout(args.attention_dropout)

    def forward(self, query_layer, key_layer, value_layer, attention_mask):
        # [sq, b, np, hn] x [sk, b, np, hn]
        sq, b, np_, hn = query_layer.shape
        sk = key_layer.size(0)
        q = query_layer.reshape(sq, b * np_, hn).transpose(0, 1)
        k = key_layer.reshape(sk, b * np_, hn).transpose(0, 1)
        scores = torch.bmm(q, k.transpose(1, 2)) * (1.0 / self.norm_factor)
        scores = scores.view(b, np_, sq, sk)
        probs = self.scale_mask_softmax(scores, attention_mask)
        if not self.sequence_parallel:
            with tp_random.get_cuda_rng_tracker().fork():
                probs = self.attention_dropout(probs)
        else:
            probs = self.attention_dropout(probs)
        v = value_layer.reshape(sk, b * np_, hn).transpose(0, 1)
        ctx = torch.bmm(probs.view(b * np_, sq, sk).to(v.dtype), v)
        ctx = ctx.view(b, np_, sq, hn).permute(2, 0, 1, 3).contiguous()
        return ctx.view(sq, b, self.hidden_size_per_partition)


class ParallelAttention(MegatronModule):
    """Self / cross attention. Self-attention uses ONE fused projection whose rows are
    grouped per KV head: [q_0 .. q_{n/g-1}, k, v] x g (checkpoint layout)."""

    def __init__(self, init_method, output_layer_init_method, layer_number, attention_type=AttnType.self_attn,
                 attn_mask_type=AttnMaskType.padding, world_size=None, args=None):
        super().__init__()
        args = args or global_vars.get_args()
        self.layer_number = max(1, layer_number)
        self.attention_type, self.attn_mask_type = attention_type, attn_mask_type
        self.params_dtype = args.params_dtype
        self.sequence_parallel = args.sequence_parallel
        self.use_flash_attn = args.use_flash_attn
        self.num_attention_heads_kv = args.num_attention_heads_kv
        self.num_attention_heads = args.num_attention_heads
        self.seq_length = args.seq_length
        self.position_embedding_type = args.position_embedding_type
        self.attention_dropout_p = args.attention_dropout
        if self.use_flash_attn:
            assert attention_type == AttnType.self_attn, "fused attention only supports self-attention"
            assert attn_mask_type == AttnMaskType.causal, "fused attention only supports causal masks"
        proj = args.kv_channels * args.num_attention_heads
        ws = world_size if world_size is not None else mpu.get_tensor_model_parallel_world_size()
        self.world_size = ws
        self.hidden_size_per_attention_head = divide(proj, args.num_attention_heads)
        self.num_attention_heads_per_partition = divide(args.num_attention_heads, ws)
        kw = _args_to_kwargs(args)
        if attention_type == AttnType.self_attn:
            kv_proj = args.kv_channels * args.num_attention_heads_kv
            assert args.num_attention_heads % args.num_attention_heads_kv == 0
            self.num_kv_heads_per_partition = divide(args.num_attention_heads_kv, ws)
            self.query_key_value = tp_layers.ColumnParallelLinear(
                args.hidden_size, proj + 2 * kv_proj, bias=args.use_bias, gather_output=False,
                init_method=init_method,
                async_tensor_model_parallel_allreduce=args.async_tensor_model_parallel_allreduce,
                world_size=ws, **kw)
        else:
            assert attention_type == AttnType.cross_attn
            self.num_kv_heads_per_partition = self.num_attention_heads_per_partition
            self.query = tp_layers.ColumnParallelLinear(
                args.hidden_size, proj, bias=args.use_bias, gather_output=False, init_method=init_method,
                async_tensor_model_parallel_allreduce=args.async_tensor_model_parallel_allreduce,
                world_size=ws, **kw)
            self.key_value = tp_layers.ColumnParallelLinear(
                args.hidden_size, 2 * proj, bias=args.use_bias, gather_output=False, init_method=init_method,
                async_tensor_model_parallel_allreduce=args.async_tensor_model_parallel_allreduce,
                world_size=ws, **kw)
        self.core_attention = CoreAttention(self.layer_number, args, self.attn_mask_type, world_size=ws)
        self.checkpoint_core_attention = args.recompute_granularity == "selective"
        self.dense = tp_layers.RowParallelLinear(
            proj, args.hidden_size, bias=args.use_bias, input_is_parallel=True,
            init_method=output_layer_init_method, skip_bias_add=True, world_size=ws, **kw)
        self.freqs_cis = None
        if self.position_embedding_type == PositionEmbeddingType.rotary:
            self._rope_cfg = (args.hidden_size // args.num_attention_heads,
                              args.max_position_embeddings, args.rope_scaling_factor)

    def _rope_table(self, device):
        if self.freqs_cis is None or self.freqs_cis.device != device:
            hd, maxp, sf = self._rope_cfg
            self.freqs_cis = precompute_freqs_cis(hd, maxp, scaling_factor=sf, device=device)
        return self.freqs_cis

    def _allocate_memory(self, inference_max_sequence_len, batch_size, device):
        return torch.empty(inference_max_sequence_len, batch_size, self.num_kv_heads_per_partition,
                           self.hidden_size_per_attention_head, dtype=self.params_dtype, device=device)

    def _checkpointed_core(self, q, k, v, mask):
        def fn(*inp):
            return self.core_attention(inp[0], inp[1], inp[2], inp[3])
        return tp_random.checkpoint(fn, False, q, k, v, mask)

    def forward(self, hidden_states, attention_mask, encoder_output=None, inference_params=None,
                position_ids=None):
        hn = self.hidden_size_per_attention_head
        np_ = self.num_attention_heads_per_partition
        if self.attention_type == AttnType.self_attn:
            mixed, _ = self.query_key_value(hidden_states)
            sq, b = mixed.shape[0], mixed.shape[1]
            g = self.num_kv_heads_per_partition
            qpg = np_ // g
            qkv = mixed.view(sq, b, g, qpg + 2, hn)
            rotary = self.position_embedding_type == PositionEmbeddingType.rotary
            fused_ok = (self.use_flash_attn and inference_params is None
                        and self.attn_mask_type == AttnMaskType.causal)
            if fused_ok:
                if rotary:
                    qkv = Fn.rope_fused_qkv(qkv, self._rope_table(qkv.device), position_ids)
                p = self.attention_dropout_p if self.training else 0.0
                if p > 0.0 and not self.sequence_parallel:
                    # inside a sequence-parallel stack the TP RNG stream is already active
                    with tp_random.get_cuda_rng_tracker().fork():
                        context = Fn.attention_fused_qkv(qkv, None, True, dropout_p=p)
                else:
                    context = Fn.attention_fused_qkv(qkv, None, True, dropout_p=p)
                return self.dense(context)
            # ---- unfused path: explicit q, k, v ----
            pos_off = 0
            if inference_params is not None:
                pos_off = inference_params.sequence_len_offset
            if rotary:
                qkv = Fn.rope_fused_qkv(qkv, self._rope_table(qkv.device), position_ids,
                                        pos_off if position_ids is None else 0)
            query = qkv[:, :, :, :qpg].reshape(sq, b, np_, hn)
            key, value = qkv[:, :, :, qpg], qkv[:, :, :, qpg + 1]
        else:
            q, _ = self.query(hidden_states)
            kv, _ = self.key_value(encoder_output)
            sq, b = q.shape[0], q.shape[1]
            query = q.view(sq, b, np_, hn)
            kv = kv.view(kv.shape[0], b, np_, 2 * hn)
            key, value = kv[..., :hn], kv[..., hn:]
            g, qpg = np_, 1

        if inference_params is not None:
            if self.layer_number not in inference_params.key_value_memory_dict:
                mem = (self._allocate_memory(inference_params.max_sequence_len,
                                             inference_params.max_batch_size, key.device),
                       self._allocate_memory(inference_params.max_sequence_len,
                                             inference_params.max_batch_size, key.device))
                inference_params.key_value_memory_dict[self.layer_number] = mem
            kmem, vmem = inference_params.key_value_memory_dict[self.layer_number]
            bs, be = inference_params.batch_size_offset, inference_params.batch_size_offset + key.size(1)
            ss, se = inference_params.sequence_len_offset, inference_params.sequence_len_offset + key.size(0)
            assert be <= kmem.size(1) and se <= kmem.size(0)
            kmem[ss:se, bs:be] = key      # keys are cached already rotated
            vmem[ss:se, bs:be] = value
            key, value = kmem[:se, bs:be], vmem[:se, bs:be]
        if g != np_:
            key = key.repeat_interleave(qpg, dim=2)
            value = value.repeat_interleave(qpg, dim=2)
        if self.checkpoint_core_attention and self.training:
            context = self._checkpointed_core(query, key, value, attention_mask)
        else:
            context = self.core_attention(query, key, value, attention_mask)
        return self.dense(context)


class ParallelTransformerLayer(MegatronModule):
    """One block: pre-LN (default), post-LN, or Falcon parallel attention + MLP."""

    def __init__(self, init_method, output_layer_init_method, layer_number, layer_type=LayerType.encoder,
                 self_attn_mask_type=AttnMaskType.padding, drop_path_rate=0.0, world_size=None,
                 hidden_dropout=0.0, args=None):
        super().__init__()
        args = args or global_vars.get_args()
        self.layer_number, self.layer_type = layer_number, layer_type
        self.apply_residual_connection_post_layernorm = args.apply_residual_connection_post_layernorm
        self.bf16 = args.bf16
        self.fp32_residual_connection = args.fp32_residual_connection
        self.parallel_attn = args.parallel_attn
        self.parallel_layernorm = args.parallel_layernorm
        self.use_rms_norm = args.use_rms_norm
        self.hidden_dropout = hidden_dropout
        self.bias_dropout_fusion = args.bias_dropout_fusion
        self.use_post_ln = args.use_post_ln
        self.sequence_parallel = args.sequence_parallel

        def make_norm():
            if args.use_rms_norm:
                return RMSNorm(args.hidden_size, eps=args.layernorm_epsilon,
                               sequence_parallel=args.sequence_parallel)
            return LayerNorm(args.hidden_size, eps=args.layernorm_epsilon,
                             no_persist_layer_norm=args.no_persist_layer_norm,
                             sequence_parallel=args.sequence_parallel)

        if args.use_post_ln:
            self.input_layernorm = torch.nn.Identity()
        else:
            self.input_layernorm = make_norm()
        self.self_attention = ParallelAttention(init_method, output_layer_init_method, layer_number,
                                                attention_type=AttnType.self_attn,
                                                attn_mask_type=self_attn_mask_type, world_size=world_size,
                                                args=args)
        self.drop_path = DropPath(drop_path_rate) if drop_path_rate > 0.0 else None
        if not self.parallel_attn:
            self.post_attention_layernorm = make_norm()
        if self.layer_type == LayerType.decoder:
            self.inter_attention = ParallelAttention(init_method, output_layer_init_method, layer_number,
                                                     attention_type=AttnType.cross_attn,
                                                     world_size=world_size, args=args)
            self.post_inter_attention_layernorm = make_norm()
        if self.parallel_layernorm:
            self.mlp_layernorm = make_norm()
        self.mlp = ParallelMLP(init_method, output_layer_init_method, args, world_size)
        self.output_layernorm = make_norm() if args.use_post_ln else torch.nn.Identity()

    def _bda(self, x, bias, residual):
        """residual + dropout(x + bias) under the right RNG stream."""
        if self.drop_path is not None:
            t = x if bias is None else x + bias
            t = F.dropout(t, p=self.hidden_dropout, training=self.training)
            return residual + self.drop_path(t)
        p = self.hidden_dropout if self.training else 0.0
        # under sequence parallelism ParallelTransformer.forward already runs inside the
        # tensor-parallel RNG stream, so shards draw different masks
        return Fn.bias_dropout_add(x, bias, residual, p, self.training)

    def forward(self, hidden_states, attention_mask, encoder_output=None, enc_dec_attn_mask=None,
                inference_params=None, position_ids=None):
        ln_out = self.input_layernorm(hidden_states)
        attn_out, attn_bias = self.self_attention(ln_out, attention_mask, inference_params=inference_params,
                                                  position_ids=position_ids)
        residual = ln_out if self.apply_residual_connection_post_layernorm else hidden_states

        if self.parallel_attn:
            # Falcon: attention and MLP read (norms of) the same input; one residual add
            mlp_in = self.mlp_layernorm(hidden_states) if self.parallel_layernorm else ln_out
            mlp_out, mlp_bias = self.mlp(mlp_in)
            both = mlp_out + attn_out
            if attn_bias is not None:
                both = both + attn_bias
            out = self._bda(both, mlp_bias, residual)
            out = self.output_layernorm(out)
            return make_viewless_tensor(out, out.requires_grad, True)

        can_fuse_norm = (self.use_rms_norm and attn_bias is None and self.drop_path is None
                         and (self.hidden_dropout == 0.0 or not self.training)
                         and not self.apply_residual_connection_post_layernorm
                         and self.layer_type == LayerType.encoder)
        if can_fuse_norm:
            # residual add + RMSNorm in one kernel: returns (norm(h), h) with h = x + residual
            ln2, ln_input = self.post_attention_layernorm(attn_out, residual)
        else:
            ln_input = self._bda(attn_out, attn_bias, residual)
            ln2 = self.post_attention_layernorm(ln_input)

        if self.layer_type == LayerType.decoder:
            inter_out, inter_bias = self.inter_attention(ln2, enc_dec_attn_mask, encoder_output=encoder_output)
            residual = ln2 if self.apply_residual_connection_post_layernorm else ln_input
            ln_input = self._bda(inter_out, inter_bias, residual)
            ln2 = self.post_inter_attention_layernorm(ln_input)

        mlp_out, mlp_bias = self.mlp(ln2)
        residual = ln2 if self.apply_residual_connection_post_layernorm else ln_input
        out = self._bda(mlp_out, mlp_bias, residual)
        out = self.output_layernorm(out)
        return make_viewless_tensor(out, out.requires_grad, True)


class NoopTransformerLayer(MegatronModule):
    """Placeholder for a stage that owns no layers (standalone embedding stage)."""

    def __init__(self, layer_number):
        super().__init__()
        self.layer_number = layer_number

    def forward(self, hidden_states, attention_mask, encoder_output=None, enc_dec_attn_mask=None,
                inference_params=None, position_ids=None):
        return hidden_states.clone()


def _get_num_layers(args, is_encoder_and_decoder_model, is_decoder=False):
    """Transformer layers owned by this pipeline rank (transformer.py:844-893 behaviour)."""
    pp = mpu.get_pipeline_model_parallel_world_size()
    # the reference leaves decoder_num_layers unset unless the flag is given (its T5 recipes
    # only pass --num_layers); an unset value means "as deep as the encoder"
    dec = args.decoder_num_layers if args.decoder_num_layers is not None else args.encoder_num_layers
    if pp > 1:
        if is_encoder_and_decoder_model:
            assert args.pipeline_model_parallel_split_rank is not None
            n_enc_ranks = (args.pipeline_model_parallel_split_rank - 1 if args.standalone_embedding_stage
                           else args.pipeline_model_parallel_split_rank)
            n_dec_ranks = args.transformer_pipeline_model_parallel_size - n_enc_ranks
            assert args.encoder_num_layers % n_enc_ranks == 0
            assert dec % n_dec_ranks == 0
            if mpu.is_pipeline_stage_before_split():
                n = 0 if args.standalone_embedding_stage and mpu.get_pipeline_model_parallel_rank() == 0 \
                    else args.encoder_num_layers // n_enc_ranks
            else:
                n = dec // n_dec_ranks
        else:
            assert args.num_layers == args.encoder_num_layers
            assert args.num_layers % args.transformer_pipeline_model_parallel_size == 0, \
                "num_layers must be divisible by transformer_pipeline_model_parallel_size"
            n = 0 if args.standalone_embedding_stage and mpu.get_pipeline_model_parallel_rank() == 0 \
                else args.num_layers // args.transformer_pipeline_model_parallel_size
    else:
        n = dec if is_decoder else args.encoder_num_layers
    return n


class ParallelTransformer(MegatronModule):
    def __init__(self, init_method, output_layer_init_method, layer_type=LayerType.encoder,
                 self_attn_mask_type=AttnMaskType.padding, post_layer_norm=True, pre_process=True,
                 post_process=True, drop_path_rate=0.0, args=None, model_type=None):
        super().__init__()
        args = args or global_vars.get_args()
        self.layer_type = layer_type
        self.model_type = model_type if model_type is not None else getattr(args, "model_type", ModelType.encoder_or_decoder)
        self.bf16 = args.bf16
        self.fp32_residual_connection = args.fp32_residual_connection
        self.post_layer_norm = post_layer_norm
        self.pre_process, self.post_process = pre_process, post_process
        self.input_tensor = None
        self.drop_path_rate = drop_path_rate
        self.transformer_impl = args.transformer_impl
        if self.transformer_impl == "transformer_engine":
            # Transformer Engine is not part of this build: its role (fused layers, FP8 linears under
            # --fp8_e4m3 / --fp8_hybrid) is played by the local layers on the sm_100a kernels (ops/fp8.py)
            from ..utils import print_rank_0
            print_rank_0("--transformer_impl transformer_engine: running the local layers on the sm_100a "
                         "kernels (FP8 linears follow --fp8_e4m3 / --fp8_hybrid)")
            self.transformer_impl = "local"
        self.recompute_granularity = args.recompute_granularity
        self.recompute_method = args.recompute_method
        self.recompute_num_layers = args.recompute_num_layers
        self.distribute_saved_activations = args.distribute_saved_activations and not args.sequence_parallel
        self.sequence_parallel = args.sequence_parallel
        self.use_post_ln = args.use_post_ln
        self.num_layers = _get_num_layers(args, self.model_type == ModelType.encoder_and_decoder,
                                          layer_type == LayerType.decoder)
        self.drop_path_rates = [r.item() for r in torch.linspace(0, self.drop_path_rate, args.num_layers)]
        n_total = max(args.num_layers, 1)
        if args.lima_dropout:
            dropouts = [args.hidden_dropout * i / max(n_total - 1, 1) for i in range(n_total)]
        else:
            dropouts = [args.hidden_dropout] * n_total

        def build_layer(layer_number):
            return ParallelTransformerLayer(
                init_method, output_layer_init_method, layer_number, layer_type=layer_type,
                self_attn_mask_type=self_attn_mask_type,
                drop_path_rate=self.drop_path_rates[min(layer_number - 1, len(self.drop_path_rates) - 1)],
                hidden_dropout=dropouts[min(layer_number - 1, n_total - 1)], args=args)

        vpp = args.virtual_pipeline_model_parallel_size
        pp_rank = mpu.get_pipeline_model_parallel_rank()
        if vpp is not None:
            assert args.num_layers % vpp == 0, "num_layers must be divisible by virtual_pipeline_model_parallel_size"
            assert self.model_type != ModelType.encoder_and_decoder
            # chunk c of stage r holds layers [c * (L / vpp) + r * n, ... + n)
            self.num_layers = self.num_layers // vpp
            offset = mpu.get_virtual_pipeline_model_parallel_rank() * (args.num_layers // vpp) + \
                pp_rank * self.num_layers
        elif self.model_type == ModelType.encoder_and_decoder and mpu.get_pipeline_model_parallel_world_size() > 1:
            if layer_type == LayerType.encoder:
                offset = pp_rank * self.num_layers
            else:
                offset = (pp_rank - args.pipeline_model_parallel_split_rank) * self.num_layers
        else:
            offset = pp_rank * self.num_layers
            if args.standalone_embedding_stage and mpu.get_pipeline_model_parallel_world_size() > 1:
                offset = max(pp_rank - 1, 0) * self.num_layers
        if self.num_layers == 0:
            self.num_layers = 1
            self.layers = torch.nn.ModuleList([NoopTransformerLayer(1)])
        else:
            self.layers = torch.nn.ModuleList([build_layer(i + 1 + offset) for i in range(self.num_layers)])
        if self.post_process and self.post_layer_norm and not self.use_post_ln:
            if args.use_rms_norm:
                self.final_layernorm = RMSNorm(args.hidden_size, eps=args.layernorm_epsilon,
                                               sequence_parallel=args.sequence_parallel)
            else:
                self.final_layernorm = LayerNorm(args.hidden_size, eps=args.layernorm_epsilon,
                                                 no_persist_layer_norm=args.no_persist_layer_norm,
                                                 sequence_parallel=args.sequence_parallel)

    def _get_layer(self, n):
        return self.layers[n]

    def _checkpointed_forward(self, hidden_states, attention_mask, encoder_output, enc_dec_attn_mask,
                              position_ids):
        def custom(start, end):
            def fwd(*inputs):
                x, mask, enc, edm, pos = inputs
                for i in range(start, end):
                    x = self._get_layer(i)(x, mask, enc, edm, position_ids=pos)
                return x
            return fwd

        if self.recompute_method == "uniform":
            l = 0
            while l < self.num_layers:
                hidden_states = tp_random.checkpoint(
                    custom(l, min(l + self.recompute_num_layers, self.num_layers)),
                    self.distribute_saved_activations, hidden_states, attention_mask, encoder_output,
                    enc_dec_attn_mask, position_ids)
                l += self.recompute_num_layers
        elif self.recompute_method == "block":
            for l in range(self.num_layers):
                if l < self.recompute_num_layers:
                    hidden_states = tp_random.checkpoint(
                        custom(l, l + 1), self.distribute_saved_activations, hidden_states, attention_mask,
                        encoder_output, enc_dec_attn_mask, position_ids)
                else:
                    hidden_states = custom(l, l + 1)(hidden_states, attention_mask, encoder_output,
                                                     enc_dec_attn_mask, position_ids)
        else:
            raise ValueError("Invalid activation recompute method.")
        return hidden_states

    def set_input_tensor(self, input_tensor):
        """Activation received from the previous pipeline stage."""
        self.input_tensor = input_tensor

    def forward(self, hidden_states, attention_mask, encoder_output=None, enc_dec_attn_mask=None,
                inference_params=None, position_ids=None):
        if inference_params:
            assert self.recompute_granularity is None, "inference does not work with activation checkpointing"
        if not self.pre_process:
            hidden_states = self.input_tensor
        hidden_states = make_viewless_tensor(hidden_states, requires_grad=True, keep_graph=True)
        rng_ctx = tp_random.get_cuda_rng_tracker().fork() if self.sequence_parallel else nullcontext()
        with rng_ctx:
            if self.recompute_granularity == "full" and self.training:
                hidden_states = self._checkpointed_forward(hidden_states, attention_mask, encoder_output,
                                                           enc_dec_attn_mask, position_ids)
            else:
                for i in range(self.num_layers):
                    hidden_states = self._get_layer(i)(
                        hidden_states, attention_mask, encoder_output=encoder_output,
                        enc_dec_attn_mask=enc_dec_attn_mask, inference_params=inference_params,
                        position_ids=position_ids)
        if self.post_process and self.post_layer_norm and not self.use_post_ln:
            hidden_states = self.final_layernorm(hidden_states)
        return hidden_states
