"""Token-overlap F1 used for the knowledge / response generation scores
(tasks/msdp/metrics.py:18-77 surface)."""
import re
from collections import Counter
from typing import List

_ARTICLES = re.compile(r"\b(a|an|the)\b")
_PUNCT = re.compile(r"[!\"#$%&()*+,-./:;<=>?@\[\]\\^`{|}~_']")


def normalize_answer(s):
    """Lower-case, drop punctuation and articles, collapse whitespace."""
    s = _PUNCT.sub(" ", s.lower())
    s = _ARTICLES.sub(" ", s)
    return " ".join(s.split())


class F1Metric:
    @staticmethod
    def _prec_recall_f1_score(pred_items, gold_items):
        common = Counter(gold_items) & Counter(pred_items)
        same = sum(common.values())
        if same == 0:
            return 0, 0, 0
        p, r = same / len(pred_items), same / len(gold_items)
        return p, r, 2 * p * r / (p + r)

    @staticmethod
    def compute_each_pair(guess: str, answer: str):
        if answer == "":
            return None, None, None
        if guess == "":
            return 0, 0, 0
        return F1Metric._prec_recall_f1_score(normalize_answer(guess).split(), normalize_answer(answer).split())

    @staticmethod
    def compute_all_pairs(guesses: List[str], answers: List[str]):
        assert len(guesses) == len(answers)
        ps, rs, fs = [], [], []
        for g, a in zip(guesses, answers):
            p, r, f = F1Metric.compute_each_pair(g, a)
            if p is None:
                continue
            ps.append(p)
            rs.append(r)
            fs.append(f)
        n = max(len(fs), 1)
        return sum(ps) / n, sum(rs) / n, sum(fs) / n
