"""Llama and Llama-2 / CodeLlama: GPTModel under the Llama argument contract (models/_family.py).
Parity: megatron/model/llama_model.py:10-43."""
from .. import global_vars
from ._family import check_family
from .gpt_model import GPTModel


class LlamaModel(GPTModel):
    def __init__(self, num_tokentypes: int = 0, parallel_output: bool = True, pre_process: bool = True,
                 post_process: bool = True, model_type=None, version: int = 2, args=None):
        if version not in (1, 2):
            raise AssertionError(f"Unknown llama version {version}")
        args = args or global_vars.get_args()
        check_family("Llama", args)
        super().__init__(num_tokentypes=num_tokentypes, parallel_output=parallel_output, pre_process=pre_process,
                         post_process=post_process, model_type=model_type, args=args)
        self.version = version
