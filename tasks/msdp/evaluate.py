"""F1 of generated knowledge / responses against references (tasks/msdp/evaluate.py:11-45 surface)."""
from epfl_megatron_b200 import get_args, print_rank_0
from msdp.metrics import F1Metric


def evaluate_f1(guess_file, answer_file):
    def read(path, strip_eos):
        out = []
        with open(path, encoding="utf-8") as f:
            for line in f:
                line = line.strip()
                if strip_eos and "<|endoftext|>" in line:
                    line = line.replace("<|endoftext|>", "")
                if not strip_eos and line == "no_passages_used":
                    line = ""
                out.append(line)
        return out
    guesses, answers = read(guess_file, True), read(answer_file, False)
    assert len(guesses) == len(answers), "lengths of guess and answer are different!"
    p, r, f1 = F1Metric.compute_all_pairs(guesses, answers)
    print_rank_0(f"Precision: {p:.4f}; recall: {r:.4f}; f1: {f1:.4f}")
    print_rank_0("done :-)")
    return p, r, f1


def main():
    args = get_args()
    evaluate_f1(args.guess_file, args.answer_file)
