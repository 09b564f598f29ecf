"""Finished-hypothesis pool for beam search (megatron/text_generation/beam_utils.py:19-64)."""
from __future__ import annotations


class BeamHypotheses:
    """Keeps the `num_beams` best finished hypotheses, scored by sum_logprobs / len**penalty."""

    def __init__(self, num_beams, length_penalty=1.0, early_stopping=False):
        self.length_penalty = length_penalty
        self.early_stopping = early_stopping
        self.num_beams = num_beams
        self.beams = []            # (score, tokens)
        self.worst_score = 1e9

    def __len__(self):
        return len(self.beams)

    def add(self, hyp, sum_logprobs, length):
        score = sum_logprobs / length ** self.length_penalty
        if len(self) < self.num_beams or score > self.worst_score:
            self.beams.append((score, hyp))
            if len(self) > self.num_beams:
                worst = min(range(len(self.beams)), key=lambda i: self.beams[i][0])
                del self.beams[worst]
            self.worst_score = min(s for s, _ in self.beams)

    def is_done(self, best_sum_logprobs, cur_len):
        """No open beam can still beat the worst kept hypothesis."""
        if len(self) < self.num_beams:
            return False
        if self.early_stopping:
            return True
        return self.worst_score >= best_sum_logprobs / cur_len ** self.length_penalty
