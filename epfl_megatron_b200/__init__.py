"""B200-native 3D-parallel LLM training framework with the capabilities and public surface
of epfLLM Megatron-LLM. See DESIGN.md; ``import megatron`` gives the reference's import
paths as a thin facade over this package."""
from .global_vars import (get_args, get_current_global_batch_size, get_num_microbatches,  # noqa: F401
                          get_signal_handler, get_tensorboard_writer, get_timers, get_tokenizer,
                          get_adlr_autoresume, get_counters, update_num_microbatches)
from .initialize import initialize_megatron  # noqa: F401
from .utils import is_last_rank, print_rank_0, print_rank_last  # noqa: F401

__version__ = "0.1.0"
