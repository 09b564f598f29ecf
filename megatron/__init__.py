"""Import-compatible facade: the reference's ``megatron.*`` module paths mapped onto
``epfl_megatron_b200``. Lets scripts written against the reference (``from megatron import
get_args``, ``from megatron.core import mpu, tensor_parallel``, ``from megatron.model import
LlamaModel`` ...) run unchanged and lets ``torch.load`` unpickle reference checkpoints whose
``args`` Namespace holds ``megatron.model.enums.PositionEmbeddingType`` members."""
import importlib
import sys
import types

_PKG = "epfl_megatron_b200"

_ALIASES = {
    "arguments": "arguments", "global_vars": "global_vars", "initialize": "initialize",
    "training": "training", "schedules": "schedules", "p2p_communication": "p2p_communication",
    "checkpointing": "checkpointing", "microbatches": "microbatches",
    "optimizer_param_scheduler": "optimizer_param_scheduler", "timers": "timers", "utils": "utils",
    "memory": "memory", "dist_signal_handler": "dist_signal_handler", "wandb_logger": "wandb_logger",
    "optimizer": "optimizer", "optimizer.optimizer": "optimizer.optimizer",
    "optimizer.distrib_optimizer": "optimizer.distrib_optimizer", "optimizer.clip_grads": "optimizer.clip_grads",
    "optimizer.grad_scaler": "optimizer.grad_scaler",
    "model": "models", "model.enums": "models.enums", "model.transformer": "models.transformer",
    "model.language_model": "models.language_model", "model.gpt_model": "models.gpt_model",
    "model.llama_model": "models.llama_model", "model.falcon_model": "models.falcon_model",
    "model.module": "models.module", "model.distributed": "models.distributed",
    "model.fused_layer_norm": "models.fused_layer_norm", "model.fused_softmax": "models.fused_softmax",
    "model.fused_bias_gelu": "models.fused_bias_gelu", "model.glu_activations": "models.glu_activations",
    "model.positional_embeddings": "models.positional_embeddings", "model.utils": "models.utils",
    "model.bert_model": "models.bert_model", "model.t5_model": "models.t5_model",
    "model.biencoder_model": "models.biencoder_model", "model.classification": "models.classification",
    "model.multiple_choice": "models.multiple_choice",
    "core.parallel_state": "parallel.state", "core.tensor_parallel": "parallel", "core.utils": "parallel.utils",
    "core.tensor_parallel.layers": "parallel.layers", "core.tensor_parallel.mappings": "parallel.mappings",
    "core.tensor_parallel.cross_entropy": "parallel.cross_entropy", "core.tensor_parallel.random": "parallel.random",
    "core.tensor_parallel.data": "parallel.data", "core.tensor_parallel.utils": "parallel.utils",
    "data": "data", "data.indexed_dataset": "data.indexed_dataset", "data.gpt_dataset": "data.gpt_dataset",
    "data.blendable_dataset": "data.blendable_dataset", "data.data_samplers": "data.data_samplers",
    "data.dataset_utils": "data.dataset_utils", "data.bert_dataset": "data.bert_dataset",
    "data.t5_dataset": "data.t5_dataset", "data.ict_dataset": "data.ict_dataset",
    "tokenizer": "tokenizer", "tokenizer.tokenizer": "tokenizer.tokenizer",
    "text_generation": "text_generation", "text_generation.api": "text_generation.api",
    "text_generation.generation": "text_generation.engine",
    "text_generation.forward_step": "text_generation.engine",
    "text_generation.sampling": "text_generation.sampling",
    "text_generation.tokenization": "text_generation.tokenization",
    "text_generation.communication": "text_generation.communication",
    "text_generation.beam_utils": "text_generation.beam_utils",
    "text_generation_server": "text_generation_server",
    "data.realm_index": "data.realm_index", "data.realm_dataset_utils": "data.realm_dataset_utils",
    "data.biencoder_dataset_utils": "data.biencoder_dataset_utils",
    "data.orqa_wiki_dataset": "data.orqa_wiki_dataset", "data.helpers": "data.helpers",
    "tokenizer.bert_tokenization": "tokenizer.bert_tokenization",
    "tokenizer.gpt2_tokenization": "tokenizer.gpt2_tokenization",
    "indexer": "indexer",
    "fp16_deprecated": "fp16_deprecated", "fp16_deprecated.loss_scaler": "fp16_deprecated.loss_scaler",
    "data.image_folder": "data.image_folder", "data.autoaugment": "data.autoaugment",
    "fused_kernels": "fused_kernels",
}


def _alias(name, target):
    full = f"megatron.{name}"
    try:
        mod = importlib.import_module(f"{_PKG}.{target}")
    except ImportError:
        return None
    sys.modules[full] = mod
    parent, _, leaf = full.rpartition(".")
    if parent in sys.modules:
        setattr(sys.modules[parent], leaf, mod)
    return mod


core = types.ModuleType("megatron.core")
core.__path__ = []
sys.modules["megatron.core"] = core
for _n, _t in _ALIASES.items():
    _alias(_n, _t)
core.mpu = sys.modules.get("megatron.core.parallel_state")
core.parallel_state = core.mpu
core.tensor_parallel = sys.modules.get("megatron.core.tensor_parallel")
core.utils = sys.modules.get("megatron.core.utils")
mpu = core.mpu

fp16_deprecated = sys.modules.get("megatron.fp16_deprecated")   # unpickling stubs for legacy checkpoints

from epfl_megatron_b200 import (get_args, get_current_global_batch_size, get_num_microbatches,  # noqa: E402,F401
                                get_signal_handler, get_tensorboard_writer, get_timers, get_tokenizer,
                                get_adlr_autoresume, get_counters, update_num_microbatches,
                                initialize_megatron, is_last_rank, print_rank_0, print_rank_last)
