"""Normalisation modules (megatron/model/fused_layer_norm.py:64-139 surface).

``RMSNorm`` and ``MixedFusedLayerNorm`` call the sm_100a kernels in ops/functional on
bf16 CUDA tensors and the PyTorch restatement otherwise. Both tag their parameters with
``sequence_parallel`` so the optimizer all-reduces their gradients over the TP group."""
import numbers

import torch
from torch.nn import init
from torch.nn.parameter import Parameter

from ..ops import functional as Fn


class MixedFusedLayerNorm(torch.nn.Module):
    def __init__(self, normalized_shape, eps=1e-5, no_persist_layer_norm=True,
                 sequence_parallel=False):
        super().__init__()
        if isinstance(normalized_shape, numbers.Integral):
            normalized_shape = (normalized_shape,)
        self.normalized_shape = torch.Size(normalized_shape)
        self.eps = eps
        self.weight = Parameter(torch.empty(*normalized_shape))
        self.bias = Parameter(torch.empty(*normalized_shape))
        self.reset_parameters()
        self.no_persist_layer_norm = no_persist_layer_norm
        self.sequence_parallel = sequence_parallel
        setattr(self.weight, "sequence_parallel", sequence_parallel)
        setattr(self.bias, "sequence_parallel", sequence_parallel)

    def reset_parameters(self):
        init.ones_(self.weight)
        init.zeros_(self.bias)

    def forward(self, x):
        return Fn.layernorm(x, self.weight, self.bias, self.eps)


LayerNorm = MixedFusedLayerNorm


class RMSNorm(torch.nn.Module):
    def __init__(self, dim, eps=1e-6, sequence_parallel=False):
        super().__init__()
        self.eps = eps
        self.weight = Parameter(torch.ones(dim))
        setattr(self.weight, "sequence_parallel", sequence_parallel)

    def forward(self, x, residual=None):
        """With ``residual`` returns (norm(x + residual), x + residual) in one kernel."""
        return Fn.rmsnorm(x, self.weight, self.eps, residual)
