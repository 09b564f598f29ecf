"""ZeRO-1 style distributed optimizer (megatron/optimizer/distrib_optimizer.py:331-700).

The implementation lives next to the other flat-buffer optimizers in ``optimizer.py``
(``DistributedOptimizer``: every data-parallel rank owns a contiguous 1/dp slice of the gradient
buffer, keeps fp32 master weights and Adam moments for that slice only, reduce-scatters the
gradients, runs the fused Adam kernel on its shard and all-gathers the updated bf16 parameters).
This module keeps the reference's import path and names.
"""
from .optimizer import DistributedOptimizer  # noqa: F401


class Range:
    """Half-open index range [start, end) (distrib_optimizer.py:15-29)."""

    def __init__(self, start, end):
        self.start, self.end = start, end
        self.size = end - start

    def normalize(self, start=0):
        return Range(start, start + self.size)

    def __str__(self):
        return "%d,%d [%d]" % (self.start, self.end, self.size)


__all__ = ["DistributedOptimizer", "Range"]
