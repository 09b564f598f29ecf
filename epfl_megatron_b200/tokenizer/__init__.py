from .tokenizer import build_tokenizer  # noqa: F401
