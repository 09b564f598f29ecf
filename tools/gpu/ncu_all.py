"""Every single-GPU kernel family of libb200kernels.so once at Llama-2 7B / Falcon shapes, for ONE
`ncu --set full -k regex:b200` capture (tools/ncu_summary.py report turns it into profiles/ncu_*.md).
The selfcheck functions are used as launchers: their fp32 PyTorch restatements run too (ncu skips
them through the kernel-name filter) and the numerics are asserted on the way."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from epfl_megatron_b200.ops import selfcheck as SC   # noqa: E402

RUNS = [
    ("rmsnorm 4096x4096 (fwd/bwd, with and without the fused residual)", lambda: SC.check_rmsnorm(4096, 4096)),
    ("layernorm 2048x8192", lambda: SC.check_layernorm(2048, 8192)),
    ("rope s4096 32 heads", lambda: SC.check_rope(4096, 1, 32, 1, 128)),
    ("glu 4096x11008", lambda: SC.check_glu(4096, 11008)),
    ("bias-gelu 2048x8192", lambda: SC.check_bias_gelu(2048, 8192)),
    ("dropout-add 4096x4096", lambda: SC.check_dropout_add(4096, 4096)),
    ("softmax b1 np8 2048x2048", lambda: SC.check_softmax(1, 8, 2048, 2048)),
    ("cross entropy 4096 x 32000", lambda: SC.check_xent(4096, 32000, 2)),
    ("adam 64M", lambda: SC.check_adam(1 << 26)),
    ("casts 16M", lambda: SC.check_casts(1 << 24)),
    ("gemm nt 4096x11008x4096 (CTA pair)", lambda: SC.check_gemm("nt", 4096, 11008, 4096, kernel=2)),
    ("gemm nn dgrad 4096x4096x11008", lambda: SC.check_gemm("nn", 4096, 4096, 11008, kernel=2)),
    ("gemm tn wgrad + fp32 accumulate 11008x4096x4096", lambda: SC.check_gemm("tn", 11008, 4096, 4096, out_mode=2, kernel=2)),
    ("gemm nt single CTA 2048x2048x1024", lambda: SC.check_gemm("nt", 2048, 2048, 1024, kernel=1)),
    ("attention fwd v2 s4096 8 heads", lambda: SC.check_attn_fwd(4096, 1, 8, 1, 128, True)),
    ("attention fwd gen1 s2048", lambda: SC.check_attn_fwd(2048, 1, 4, 1, 128, True, entry="b200_attn_fwd")),
    ("attention bwd s4096 8 heads", lambda: SC.check_attn_bwd(4096, 1, 8, 1, 128, True)),
    ("attention bwd GQA s2048 2x4", lambda: SC.check_attn_bwd(2048, 1, 2, 4, 128, True)),
    ("attention bwd v2 s2048", lambda: SC.check_attn_bwd(2048, 1, 4, 1, 128, True, entry="b200_attn_bwd_v2")),
]

if __name__ == "__main__":
    only = sys.argv[1:]
    for name, fn in RUNS:
        if only and not any(o in name for o in only):
            continue
        try:
            fn()
            torch.cuda.synchronize()
            print("ok  ", name, flush=True)
        except Exception as e:  # noqa: BLE001
            print("FAIL", name, f"{type(e).__name__}: {e}"[:300], flush=True)
