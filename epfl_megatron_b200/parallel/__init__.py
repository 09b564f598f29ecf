"""Tensor-parallel toolkit (megatron/core/tensor_parallel/__init__.py:1-65 surface)."""
from .cross_entropy import vocab_parallel_cross_entropy  # noqa: F401
from .data import broadcast_data  # noqa: F401
from .layers import (ColumnParallelLinear, RowParallelLinear, VocabParallelEmbedding,  # noqa: F401
                     copy_tensor_model_parallel_attributes, linear_with_grad_accumulation_and_async_allreduce,
                     param_is_not_tensor_parallel_duplicate, set_defaults_if_not_set_tensor_model_parallel_attributes,
                     set_tensor_model_parallel_attributes)
from .mappings import (copy_to_tensor_model_parallel_region, gather_from_sequence_parallel_region,  # noqa: F401
                       gather_from_tensor_model_parallel_region, reduce_from_tensor_model_parallel_region,
                       reduce_scatter_to_sequence_parallel_region, scatter_to_sequence_parallel_region,
                       scatter_to_tensor_model_parallel_region)
from .random import checkpoint, get_cuda_rng_tracker, model_parallel_cuda_manual_seed  # noqa: F401
from .utils import (gather_split_1d_tensor, split_tensor_along_last_dim,  # noqa: F401
                    split_tensor_into_1d_equal_chunks, VocabUtility, divide)
