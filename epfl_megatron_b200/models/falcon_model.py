"""Falcon: GPTModel under the Falcon argument contract (models/_family.py): rotary, parallel
attention + MLP block, GeLU, no biases, grouped / multi-query KV heads.
Parity: megatron/model/falcon_model.py:10-41."""
from .. import global_vars
from ._family import check_family
from .gpt_model import GPTModel


class FalconModel(GPTModel):
    def __init__(self, num_tokentypes: int = 0, parallel_output: bool = True, pre_process: bool = True,
                 post_process: bool = True, model_type=None, args=None):
        args = args or global_vars.get_args()
        check_family("Falcon", args)
        super().__init__(num_tokentypes=num_tokentypes, parallel_output=parallel_output, pre_process=pre_process,
                         post_process=post_process, model_type=model_type, args=args)
