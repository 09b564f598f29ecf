"""Text generation on top of the training model classes (megatron/text_generation/ surface:
api.py:19-201). KV cache, sampling, beam search and the pipeline plumbing live in engine.py."""
from .api import (beam_search, beam_search_and_post_process, generate,  # noqa: F401
                  generate_and_post_process)
