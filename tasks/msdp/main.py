"""Multi-stage dialogue prompting entry (tasks/msdp/main.py:14-64 surface):
--task MSDP-PROMPT (generate knowledge / responses by prompting) or MSDP-EVAL-F1."""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir, os.path.pardir)))
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))

from epfl_megatron_b200 import get_args  # noqa: E402
from epfl_megatron_b200.initialize import initialize_megatron  # noqa: E402


def get_tasks_args(parser):
    g = parser.add_argument_group(title="tasks")
    g.add_argument("--task", type=str, required=True)
    g.add_argument("--sample_input_file", type=str, default=None, help="file with the test samples")
    g.add_argument("--sample_output_file", type=str, default=None, help="where the generations go")
    g.add_argument("--prompt_file", type=str, default=None)
    g.add_argument("--prompt_type", type=str, default=None, choices=["knowledge", "response"])
    g.add_argument("--num_prompt_examples", type=int, default=10)
    g.add_argument("--guess_file", type=str, default=None)
    g.add_argument("--answer_file", type=str, default=None)
    g.add_argument("--out_seq_length", type=int, default=100, help="max tokens to generate")
    g.add_argument("--api_prompt", default=False, action="store_true", help="prompt a served model over REST")
    g.add_argument("--megatron_api_url", type=str, default=None)
    g.add_argument("--model_name", default="gpt", choices=["gpt", "llama", "llama2", "codellama", "falcon"])
    return parser


if __name__ == "__main__":
    initialize_megatron(extra_args_provider=get_tasks_args)
    args = get_args()
    if args.num_layers_per_virtual_pipeline_stage is not None:
        print("Interleaved pipeline schedule is not yet supported for downstream tasks.")
        sys.exit()
    if args.task == "MSDP-PROMPT":
        from msdp.prompt import main
    elif args.task == "MSDP-EVAL-F1":
        from msdp.evaluate import main
    else:
        raise NotImplementedError(f"Task {args.task} is not implemented.")
    main()
