"""Open-retrieval QA evaluation entry (tasks/orqa/evaluate_orqa.py:9-38 surface)."""
from epfl_megatron_b200 import get_args, print_rank_0
from orqa.evaluate_utils import ORQAEvaluator


def main():
    args = get_args()
    evaluator = ORQAEvaluator()
    if args.qa_data_dev is not None:
        evaluator.evaluate(args.qa_data_dev, "DEV")
    if args.qa_data_test is not None:
        evaluator.evaluate(args.qa_data_test, "TEST")
    print_rank_0("done :-)")
