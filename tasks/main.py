"""Downstream tasks dispatcher (tasks/main.py:14-96 surface).

    python tasks/main.py --task {MNLI,QQP,RACE,LAMBADA,WIKITEXT103,ICT-ZEROSHOT-NQ,RETRIEVER-EVAL,RET-FINETUNE-NQ} ...
"""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))
sys.path.insert(0, os.path.abspath(os.path.dirname(__file__)))

from epfl_megatron_b200 import get_args  # noqa: E402
from epfl_megatron_b200.initialize import initialize_megatron  # noqa: E402


def get_tasks_args(parser):
    g = parser.add_argument_group(title="tasks")
    g.add_argument("--task", type=str, required=True, help="Task name.")
    g.add_argument("--epochs", type=int, default=None, help="finetuning epochs; 0 = evaluation only")
    g.add_argument("--pretrained_checkpoint", type=str, default=None)
    g.add_argument("--keep_last", action="store_true", help="keep the last (maybe incomplete) batch")
    g.add_argument("--train_data", nargs="+", default=None)
    g.add_argument("--valid_data", nargs="*", default=None)
    g.add_argument("--overlapping_eval", type=int, default=32, help="sliding window for overlapping evaluation")
    g.add_argument("--strict_lambada", action="store_true", help="whole-last-word formulation of LAMBADA")
    g.add_argument("--qa_data_dev", type=str, default=None)
    g.add_argument("--qa_data_test", type=str, default=None)
    g.add_argument("--faiss_use_gpu", action="store_true")
    g.add_argument("--faiss_match", type=str, default="string", choices=["regex", "string"])
    g.add_argument("--faiss_topk_retrievals", type=int, default=100)
    g.add_argument("--eval_micro_batch_size", type=int, default=None)
    g.add_argument("--train_with_neg", action="store_true")
    g.add_argument("--train_hard_neg", type=int, default=0)
    g.add_argument("--val_av_rank_hard_neg", type=int, default=30)
    g.add_argument("--val_av_rank_other_neg", type=int, default=30)
    return parser


def dispatch(task):
    if task == "RACE":
        from race.finetune import main
    elif task in ("MNLI", "QQP"):
        from glue.finetune import main
    elif task in ("LAMBADA", "WIKITEXT103"):
        from zeroshot_gpt.evaluate import main
    elif task in ("ICT-ZEROSHOT-NQ", "RETRIEVER-EVAL"):
        from orqa.evaluate_orqa import main
    elif task == "RET-FINETUNE-NQ":
        from orqa.supervised.finetune import main
    else:
        raise NotImplementedError(f"Task {task} is not implemented.")
    return main


if __name__ == "__main__":
    initialize_megatron(extra_args_provider=get_tasks_args)
    args = get_args()
    if args.num_layers_per_virtual_pipeline_stage is not None:
        print("Interleaved pipeline schedule is not yet supported for downstream tasks.")
        sys.exit()
    dispatch(args.task)()
