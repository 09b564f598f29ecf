"""Text detokenisers that undo dataset-specific tokenisation before perplexity is computed
(tasks/zeroshot_gpt/detokenizer.py:8-67 surface)."""
import re


def ptb_detokenizer(s):
    for a, b in ((" '", "'"), (" \n", "\n"), ("\n ", "\n"), (" n't", "n't"), (" N ", "1 "), ("$ 1", "$1"),
                 ("# 1", "#1")):
        s = s.replace(a, b)
    return s


def wikitext_detokenizer(s):
    s = s.replace("s '", "s'")
    s = re.sub(r"/' [0-9]/", r"/'[0-9]/", s)
    for a, b in ((" @-@ ", "-"), (" @,@ ", ","), (" @.@ ", "."), (" : ", ": "), (" ; ", "; "), (" . ", ". "),
                 (" ! ", "! "), (" ? ", "? "), (" , ", ", ")):
        s = s.replace(a, b)
    s = re.sub(r"\(\s*([^\)]*?)\s*\)", r"(\1)", s)
    s = re.sub(r"\[\s*([^\]]*?)\s*\]", r"[\1]", s)
    s = re.sub(r"{\s*([^}]*?)\s*}", r"{\1}", s)
    s = re.sub(r"\"\s*([^\"]*?)\s*\"", r'"\1"', s)
    s = re.sub(r"'\s*([^']*?)\s*'", r"'\1'", s)
    for a, b in (("= = = =", "===="), ("= = =", "==="), ("= =", "=="), (" " + chr(176) + " ", chr(176)),
                 (" \n", "\n"), ("\n ", "\n"), (" N ", " 1 "), (" 's", "'s")):
        s = s.replace(a, b)
    return s


def lambada_detokenizer(s):
    return s


_DETOKENIZERS = {"ptb": ptb_detokenizer, "wiki": wikitext_detokenizer, "lambada": lambada_detokenizer}


def get_detokenizer(path):
    for key, fn in _DETOKENIZERS.items():
        if key in path:
            return fn
    return None
