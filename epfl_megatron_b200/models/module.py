"""Module base classes (megatron/model/module.py:24-202 surface): ``MegatronModule`` with
the tied-embedding plumbing across the first and last pipeline stage, and
``Float16Module`` which holds a half/bf16 copy of the model and converts at the
pipeline ends."""
from __future__ import annotations

import torch
from torch.autograd import Variable
from torch.nn.parameter import Parameter

from .. import global_vars
from ..parallel import state as mpu
from ..parallel import layers as tp_layers

_FLOAT_TYPES = (torch.FloatTensor, torch.cuda.FloatTensor) if torch.cuda.is_available() else (torch.FloatTensor,)


def param_is_not_shared(param):
    return not hasattr(param, "shared") or not param.shared


class MegatronModule(torch.nn.Module):
    def __init__(self, share_word_embeddings=True):
        super().__init__()
        self.share_word_embeddings = share_word_embeddings

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        return self.state_dict(prefix=prefix, keep_vars=keep_vars)

    def word_embeddings_weight(self):
        if self.pre_process:
            return self.language_model.embedding.word_embeddings.weight
        if not self.share_word_embeddings:
            raise Exception("word_embeddings_weight() called for last stage, but share_word_embeddings is false")
        return self.word_embeddings.weight

    def initialize_word_embeddings(self, init_method_normal, args):
        """With tied embeddings and pp > 1 the last stage keeps a second, zero-initialised
        copy that is all-reduced with the first stage's weights (and grads each step)."""
        if not self.share_word_embeddings:
            raise Exception("initialize_word_embeddings() was called but share_word_embeddings is false")
        if args.pipeline_model_parallel_size == 1:
            return
        if mpu.is_pipeline_last_stage() and not self.pre_process:
            assert not mpu.is_pipeline_first_stage()
            self._word_embeddings_for_head_key = "word_embeddings_for_head"
            self.word_embeddings = tp_layers.VocabParallelEmbedding(
                args.padded_vocab_size, args.hidden_size,
                init_method=init_method_normal(args.init_method_std),
                params_dtype=args.params_dtype,
                use_cpu_initialization=args.use_cpu_initialization,
                perform_initialization=args.perform_initialization)
            self.word_embeddings.weight.data.fill_(0)
            self.word_embeddings.weight.shared = True
        if not mpu.is_pipeline_first_stage(ignore_virtual=True) and self.pre_process:
            self.language_model.embedding.zero_parameters()
        if not torch.distributed.is_initialized():
            return
        if mpu.is_rank_in_embedding_group():
            torch.distributed.all_reduce(self.word_embeddings_weight().data, group=mpu.get_embedding_group())
        if mpu.is_rank_in_position_embedding_group() and \
                args.pipeline_model_parallel_split_rank is not None:
            if torch.cuda.is_available() and args.distributed_backend != "gloo":
                self.language_model.embedding.cuda()
            pos = self.language_model.embedding.position_embeddings
            torch.distributed.all_reduce(pos.weight.data, group=mpu.get_position_embedding_group())


def conversion_helper(val, conversion):
    if not isinstance(val, (tuple, list)):
        return conversion(val)
    out = [conversion_helper(v, conversion) for v in val]
    return tuple(out) if isinstance(val, tuple) else out


def fp32_to_float16(val, float16_convertor):
    def conv(v):
        if isinstance(v, (Parameter, Variable)):
            t = v.data
        else:
            t = v
        if isinstance(t, torch.Tensor) and t.dtype == torch.float32:
            return float16_convertor(v)
        return v
    return conversion_helper(val, conv)


def float16_to_fp32(val):
    def conv(v):
        t = v.data if isinstance(v, (Parameter, Variable)) else v
        if isinstance(t, torch.Tensor) and t.dtype in (torch.bfloat16, torch.half):
            return v.float()
        return v
    return conversion_helper(val, conv)


class Float16Module(MegatronModule):
    def __init__(self, module, args):
        super().__init__()
        if args.fp16:
            self.add_module("module", module.half())
            self._conv = lambda v: v.half()
        elif args.bf16:
            self.add_module("module", module.bfloat16())
            self._conv = lambda v: v.bfloat16()
        else:
            raise Exception("should not be here")

    def set_input_tensor(self, input_tensor):
        return self.module.set_input_tensor(input_tensor)

    def forward(self, *inputs, **kwargs):
        if mpu.is_pipeline_first_stage():
            inputs = fp32_to_float16(inputs, self._conv)
        outputs = self.module(*inputs, **kwargs)
        if mpu.is_pipeline_last_stage():
            outputs = float16_to_fp32(outputs)
        return outputs

    def state_dict(self, prefix="", keep_vars=False):
        return self.module.state_dict(prefix=prefix, keep_vars=keep_vars)

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        return self.module.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)

    def load_state_dict(self, state_dict, strict=True):
        self.module.load_state_dict(state_dict, strict=strict)
