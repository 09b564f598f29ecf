"""Scale + mask + softmax for the non-flash attention path
(megatron/model/fused_softmax.py:102-213 surface). The sm_100a kernel has no sequence
length cap (the reference kernels stop at 2048 / 4096 keys)."""
import torch

from ..ops import functional as Fn
from .enums import AttnMaskType


class FusedScaleMaskSoftmax(torch.nn.Module):
    def __init__(self, input_in_fp16, input_in_bf16, attn_mask_type, scaled_masked_softmax_fusion,
                 mask_func, softmax_in_fp32, scale):
        super().__init__()
        self.input_in_fp16, self.input_in_bf16 = input_in_fp16, input_in_bf16
        assert not (input_in_fp16 and input_in_bf16), "both fp16 and bf16 flags cannot be active at the same time."
        self.input_in_float16 = input_in_fp16 or input_in_bf16
        self.attn_mask_type = attn_mask_type
        self.scaled_masked_softmax_fusion = scaled_masked_softmax_fusion
        self.mask_func = mask_func
        self.softmax_in_fp32 = softmax_in_fp32
        self.scale = scale
        assert self.scale is None or softmax_in_fp32, "softmax should be in fp32 when scaled"

    def is_kernel_available(self, mask, b, np, sq, sk):
        return self.scaled_masked_softmax_fusion and self.input_in_bf16 and sk % 8 == 0 and \
            Fn.kernels_enabled()

    def forward(self, input, mask):
        assert input.dim() == 4
        b, np_, sq, sk = input.shape
        scale = self.scale if self.scale is not None else 1.0
        if self.is_kernel_available(mask, b, np_, sq, sk) and input.is_cuda:
            if self.attn_mask_type == AttnMaskType.causal and sq == sk:
                return Fn.scaled_softmax(input, scale, None, causal=True)
            return Fn.scaled_softmax(input, scale, mask, causal=False)
        return self.forward_torch_softmax(input, mask)

    def forward_torch_softmax(self, input, mask):
        if self.input_in_float16 and self.softmax_in_fp32:
            input = input.float()
        if self.scale is not None:
            input = input * self.scale
        masked = self.mask_func(input, mask) if mask is not None else input
        probs = torch.nn.Softmax(dim=-1)(masked)
        if self.input_in_float16 and self.softmax_in_fp32:
            probs = probs.half() if self.input_in_fp16 else probs.bfloat16()
        return probs
