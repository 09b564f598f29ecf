"""Answer-presence scoring of retrieved passages (tasks/orqa/unsupervised/qa_utils.py:27-177
surface): a passage is a hit when it contains one of the gold answers (token-level string match
after normalisation, or a regular expression)."""
import collections
import logging
import string
import unicodedata
from functools import partial
from multiprocessing import Pool as ProcessPool

import regex as re

from orqa.unsupervised.tokenizers import SimpleTokenizer

logger = logging.getLogger(__name__)
QAMatchStats = collections.namedtuple("QAMatchStats", ["top_k_hits", "questions_doc_hits"])


def calculate_matches(all_docs, answers, closest_docs, workers_num, match_type):
    """all_docs: {doc id -> (text, title)}; closest_docs: per question (ids, scores).
    top_k_hits[k] = number of questions answered by one of their first k+1 passages."""
    global dpr_all_documents
    dpr_all_documents = all_docs
    tok = SimpleTokenizer()
    fn = partial(check_answer, match_type=match_type, tokenizer=tok)
    pairs = list(zip(answers, closest_docs))
    if workers_num and workers_num > 1:
        with ProcessPool(processes=workers_num) as pool:
            scores = pool.map(fn, pairs)
    else:
        scores = [fn(p) for p in pairs]
    n_docs = len(closest_docs[0][0])
    top_k_hits = [0] * n_docs
    for hits in scores:
        best = next((i for i, x in enumerate(hits) if x), None)
        if best is not None:
            for i in range(best, n_docs):
                top_k_hits[i] += 1
    return QAMatchStats(top_k_hits, scores)


def check_answer(questions_answers_docs, tokenizer, match_type):
    answers, (doc_ids, _scores) = questions_answers_docs
    global dpr_all_documents
    hits = []
    for doc_id in doc_ids:
        doc = dpr_all_documents.get(doc_id)
        text = doc[0] if doc is not None else None
        if text is None:
            logger.warning("no doc in db")
            hits.append(False)
            continue
        hits.append(has_answer(answers, text, tokenizer, match_type))
    return hits


def has_answer(answers, text, tokenizer, match_type):
    text = _normalize(text)
    if match_type == "string":
        words = tokenizer.tokenize(text).words(uncased=True)
        for ans in answers:
            a = tokenizer.tokenize(_normalize(ans)).words(uncased=True)
            for i in range(0, len(words) - len(a) + 1):
                if a == words[i:i + len(a)]:
                    return True
    elif match_type == "regex":
        for ans in answers:
            if regex_match(text, _normalize(ans)):
                return True
    return False


def regex_match(text, pattern):
    try:
        pat = re.compile(pattern, flags=re.IGNORECASE + re.UNICODE + re.MULTILINE)
    except BaseException:   # noqa: BLE001 - an invalid pattern never matches
        return False
    return pat.search(text) is not None


def exact_match_score(prediction, ground_truth):
    return _normalize_answer(prediction) == _normalize_answer(ground_truth)


def _normalize_answer(s):
    s = "".join(ch for ch in s.lower() if ch not in set(string.punctuation))
    s = re.sub(r"\b(a|an|the)\b", " ", s)
    return " ".join(s.split())


def _normalize(text):
    return unicodedata.normalize("NFD", text)
