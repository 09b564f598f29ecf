"""Start the generation REST server (tools/run_text_generation_server.py:49-84 surface).

    torchrun --nproc-per-node <tp*pp> tools/run_text_generation_server.py --model_name llama2 \
        --load <ckpt> --use_checkpoint_args --tokenizer_type SentencePieceTokenizer --vocab_file tokenizer.model \
        --tensor_model_parallel_size T --pipeline_model_parallel_size P --bf16 [--port 5000]

Unlike the reference (which only builds GPTModel, run_text_generation_server.py:24-32) the model
class follows ``--model_name``, so Llama / Falcon checkpoints can be served.
"""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))

import torch.distributed as dist

from epfl_megatron_b200 import get_args, print_rank_0
from epfl_megatron_b200.checkpointing import load_checkpoint
from epfl_megatron_b200.initialize import initialize_megatron
from epfl_megatron_b200.models import FalconModel, GPTModel, LlamaModel
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.text_generation_server import MegatronServer, serve_worker_loop
from epfl_megatron_b200.training import get_model


def model_provider(pre_process=True, post_process=True):
    args = get_args()
    print_rank_0(f"building {args.model_name} model for generation ...")
    cls = {"gpt": GPTModel, "falcon": FalconModel}.get(args.model_name, LlamaModel)
    kw = dict(num_tokentypes=0, parallel_output=False, pre_process=pre_process, post_process=post_process)
    if cls is LlamaModel:
        kw["version"] = 2 if args.model_name in ("llama2", "codellama") else 1
    return cls(**kw)


def add_text_generate_args(parser):
    g = parser.add_argument_group(title="text generation")
    g.add_argument("--port", type=int, default=5000, help="port of the REST server")
    g.add_argument("--host", type=str, default="0.0.0.0")
    g.add_argument("--model_name", default="gpt",
                   choices=["gpt", "llama", "llama2", "codellama", "falcon"])
    g.add_argument("--model_type", default="encoder_or_decoder")
    return parser


def main():
    initialize_megatron(extra_args_provider=add_text_generate_args,
                        args_defaults={"tokenizer_type": "GPT2BPETokenizer", "no_load_rng": True,
                                       "no_load_optim": True})
    args = get_args()
    if args.num_layers_per_virtual_pipeline_stage is not None:
        raise SystemExit("interleaved pipeline schedules are not supported for text generation")
    model = get_model(model_provider, wrap_with_ddp=False)
    if args.load is not None:
        load_checkpoint(model, None, None)
    assert len(model) == 1
    model = model[0]
    model.eval()
    if mpu.is_pipeline_first_stage() and mpu.get_tensor_model_parallel_rank() == 0 \
            and (not dist.is_initialized() or dist.get_rank() == 0):
        MegatronServer(model).run(args.host, port=args.port)
    else:
        serve_worker_loop(model)


if __name__ == "__main__":
    main()
