"""Token sampling: greedy, top-k, nucleus (megatron/text_generation/sampling.py:14-93)."""
from __future__ import annotations

import torch


def modify_logits_for_top_k_filtering(logits, top_k):
    """Keep the k largest logits per row, set the rest to -inf (in place)."""
    kth = torch.topk(logits, top_k)[0][..., -1, None]
    logits.masked_fill_(logits < kth, float("-inf"))


def modify_logits_for_top_p_filtering(logits, top_p):
    """Keep the smallest prefix of the sorted distribution whose mass exceeds top_p."""
    sorted_logits, sorted_idx = torch.sort(logits, descending=True)
    cum = sorted_logits.softmax(dim=-1).cumsum(dim=-1)
    drop = cum > top_p
    # shift right so the first token that crosses the threshold is kept
    drop[..., 1:] = drop[..., :-1].clone()
    drop[..., 0] = False
    drop = drop.scatter(1, sorted_idx, drop)
    logits.masked_fill_(drop, float("-inf"))


def sample(logits, top_k=0, top_p=0.0, temperature=1.0, vocab_size=None):
    """logits [b, v] -> token ids [b]. top_k == 1 is greedy; top_k and top_p are exclusive."""
    assert logits.dim() == 2, "expected [batch, vocab] logits"
    assert logits.dtype == torch.float32, "sample in fp32"
    if top_k == 1:
        assert top_p == 0.0, "greedy sampling takes no top-p"
        samples = torch.argmax(logits, dim=-1)
    else:
        logits = logits.clone()
        if temperature != 1.0:
            logits.div_(temperature)
        if top_k > 1:
            assert top_p == 0.0, "top-k and top-p are exclusive"
            assert top_k <= logits.size(1), "top-k larger than the vocabulary"
            if vocab_size:
                assert top_k < vocab_size, "top-k larger than the tokenizer vocabulary"
            modify_logits_for_top_k_filtering(logits, top_k)
        elif top_p > 0.0:
            assert top_p <= 1.0, "top-p must be in (0, 1]"
            modify_logits_for_top_p_filtering(logits, top_p)
        samples = torch.multinomial(logits.softmax(dim=-1), num_samples=1).view(-1)
    if vocab_size:
        # padded vocabulary entries can never be produced
        samples = torch.clamp(samples, min=0, max=vocab_size - 1)
    return samples
