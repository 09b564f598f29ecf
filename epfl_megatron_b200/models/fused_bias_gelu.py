"""tanh-GeLU(bias + y) (megatron/model/fused_bias_gelu.py:14-43 surface)."""
from ..ops import functional as Fn


def bias_gelu_impl(x, bias):
    return Fn.bias_gelu(x, bias, tanh_approx=True)
