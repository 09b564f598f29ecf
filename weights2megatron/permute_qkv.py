"""RoPE row-order conversion between HF 'half-rotation' and Meta/Megatron 'interleaved
pairs' for the q and k rows of a fused QKV weight, and a migration tool for checkpoints
that were converted without it (reference: weights2megatron/permute_qkv.py:12-72)."""
import os
import re
import sys
from argparse import ArgumentParser
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from epfl_megatron_b200.utils.ckpt_convert import permute_qkv  # noqa: E402,F401


def update_checkpoint(input_dir: Path, output_dir: Path, overwrite_ok: bool = False):
    """Rewrite every ``query_key_value.weight`` of a Megatron checkpoint tree."""
    with open(input_dir / "latest_checkpointed_iteration.txt") as f:
        it = f.read().strip()
    sub = "release" if it == "release" else f"iter_{int(it):07d}"
    for fname in sorted((input_dir / sub).rglob("*.pt")):
        out = output_dir / fname.relative_to(input_dir)
        if out.exists() and not overwrite_ok:
            raise FileExistsError(f"{out} exists; pass --overwrite_ok")
        print("processing", fname)
        sd = torch.load(fname, map_location="cpu", weights_only=False)
        a = sd["args"]
        lm = sd["model"]["language_model"]
        enc = lm.get("encoder", lm.get("transformer"))
        tp = getattr(a, "tensor_model_parallel_size", 1)
        for k in list(enc):
            if re.match(r"^layers\.[0-9]+\.(self_)?attention\.query_key_value\.weight$", k):
                # a tensor-parallel shard holds whole KV groups: local head counts, same head size
                heads, kv = a.num_attention_heads // tp, a.num_attention_heads_kv // tp
                enc[k] = permute_qkv(enc[k], (a.hidden_size // a.num_attention_heads) * heads, heads, kv)
        out.parent.mkdir(parents=True, exist_ok=True)
        torch.save(sd, out)
    (output_dir / "latest_checkpointed_iteration.txt").write_text(it)


if __name__ == "__main__":
    p = ArgumentParser(description="Fix the RoPE row order of the QKV weights of a Megatron checkpoint")
    p.add_argument("--input_dir", type=Path, required=True)
    p.add_argument("--output_dir", type=Path, required=True)
    p.add_argument("--overwrite_ok", action="store_true")
    a = p.parse_args()
    update_checkpoint(a.input_dir, a.output_dir, a.overwrite_ok)
