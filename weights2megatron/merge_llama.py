"""Bring Llama weights into Meta's un-sharded naming: merge Meta ``consolidated.NN.pth``
shards, or rename a HF ``LlamaForCausalLM`` state dict (reference:
weights2megatron/merge_llama.py:21-109)."""
import os
import re
from collections import OrderedDict
from pathlib import Path

import torch

# dimension along which Meta's model-parallel shards are concatenated (None = replicated)
_META_CAT_DIM = {"w1": 0, "w2": -1, "w3": 0, "wo": -1, "wq": 0, "wk": 0, "wv": 0, "output": 0,
                 "tok_embeddings": -1, "ffn_norm": None, "attention_norm": None, "norm": None,
                 "rope": None}


def _short(key):
    parts = key.split(".")
    return parts[-2] if len(parts) > 1 else parts[0]


def merge_meta_llama(size: int, root_dir: Path):
    paths = sorted(p for p in Path(root_dir).iterdir() if re.match(r"^consolidated\.[0-9]+\.pth$", p.name))
    assert paths, f"no consolidated.*.pth under {root_dir}"
    shards = [torch.load(p, map_location="cpu", weights_only=False) for p in paths]
    out = OrderedDict()
    for key in shards[0]:
        dim = _META_CAT_DIM.get(_short(key), None)
        if dim is None:
            out[key] = shards[0][key]
        else:
            out[key] = torch.cat([s[key] for s in shards], dim=dim)
    return out


def merge_hf_llama(size: int, version: int, cache_dir=None, model_path=None):
    """HF checkpoint -> Meta key names (values untouched: q/k rows stay in HF RoPE order)."""
    from transformers import LlamaForCausalLM
    assert model_path is not None, "offline: pass a local HF model directory"
    model = LlamaForCausalLM.from_pretrained(model_path, cache_dir=cache_dir, torch_dtype=torch.float32)
    return hf_state_to_meta(model.state_dict())


def hf_state_to_meta(sd):
    ren = [(r"^model\.embed_tokens\.weight$", "tok_embeddings.weight"),
           (r"^model\.norm\.weight$", "norm.weight"), (r"^lm_head\.weight$", "output.weight"),
           (r"^model\.layers\.(\d+)\.self_attn\.q_proj\.weight$", r"layers.\1.attention.wq.weight"),
           (r"^model\.layers\.(\d+)\.self_attn\.k_proj\.weight$", r"layers.\1.attention.wk.weight"),
           (r"^model\.layers\.(\d+)\.self_attn\.v_proj\.weight$", r"layers.\1.attention.wv.weight"),
           (r"^model\.layers\.(\d+)\.self_attn\.o_proj\.weight$", r"layers.\1.attention.wo.weight"),
           (r"^model\.layers\.(\d+)\.mlp\.gate_proj\.weight$", r"layers.\1.feed_forward.w1.weight"),
           (r"^model\.layers\.(\d+)\.mlp\.down_proj\.weight$", r"layers.\1.feed_forward.w2.weight"),
           (r"^model\.layers\.(\d+)\.mlp\.up_proj\.weight$", r"layers.\1.feed_forward.w3.weight"),
           (r"^model\.layers\.(\d+)\.input_layernorm\.weight$", r"layers.\1.attention_norm.weight"),
           (r"^model\.layers\.(\d+)\.post_attention_layernorm\.weight$", r"layers.\1.ffn_norm.weight")]
    out = OrderedDict()
    for k, v in sd.items():
        if "rotary_emb" in k:
            continue
        for pat, rep in ren:
            if re.match(pat, k):
                out[re.sub(pat, rep, k)] = v
                break
        else:
            raise KeyError(f"unexpected HF key {k}")
    return out


def merge_llama(size: int, version: int, root_dir=None, model_path=None):
    if root_dir is not None and any(re.match(r"^consolidated\.[0-9]+\.pth$", f) for f in os.listdir(root_dir)):
        return merge_meta_llama(size, Path(root_dir)), "meta"
    print(f"Weights at {root_dir} do not look like a meta checkpoint, assuming huggingface format")
    return merge_hf_llama(size, version, root_dir, model_path or root_dir), "hf"
