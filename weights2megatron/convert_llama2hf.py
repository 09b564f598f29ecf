"""Meta Llama-1 weights (consolidated.NN.pth + params.json) -> Hugging Face LlamaForCausalLM
directory (the job of weights2megatron/convert_llama2hf.py:77-323: building a verification
baseline). Shards are merged with ``merge_llama.merge_meta_llama`` and Q/K rows are permuted
from Meta's interleaved-pair RoPE order to HF's rotate-half order.

    python weights2megatron/convert_llama2hf.py --input_dir /path/to/llama --model_size 7 --output_dir out
"""
import argparse
import json
import os
import shutil
from pathlib import Path

import torch

NUM_SHARDS = {7: 1, 13: 2, 30: 4, 34: 4, 65: 8, 70: 8}


def compute_intermediate_size(n, ffn_dim_multiplier=1.0, multiple_of=256):
    return multiple_of * ((int(ffn_dim_multiplier * int(8 * n / 3)) + multiple_of - 1) // multiple_of)


def to_hf_rope_order(w, n_heads, dim_out, dim_in):
    """[(head, pair, 2), in] -> [(head, 2, pair), in]"""
    return w.view(n_heads, dim_out // n_heads // 2, 2, dim_in).transpose(1, 2).reshape(dim_out, dim_in)


def write_model(model_path, input_base_path, model_size, safe_serialization=True):
    from transformers import LlamaConfig, LlamaForCausalLM
    from merge_llama import merge_meta_llama
    input_base_path = Path(input_base_path)
    with open(input_base_path / "params.json") as f:
        params = json.load(f)
    n_layers, n_heads, dim = params["n_layers"], params["n_heads"], params["dim"]
    n_kv = params.get("n_kv_heads", n_heads)
    hd = dim // n_heads
    sd = merge_meta_llama(model_size, input_base_path)
    out = {"model.embed_tokens.weight": sd["tok_embeddings.weight"], "model.norm.weight": sd["norm.weight"],
           "lm_head.weight": sd["output.weight"]}
    for i in range(n_layers):
        p, q = f"layers.{i}.", f"model.layers.{i}."
        out[q + "self_attn.q_proj.weight"] = to_hf_rope_order(sd[p + "attention.wq.weight"], n_heads, dim, dim)
        out[q + "self_attn.k_proj.weight"] = to_hf_rope_order(sd[p + "attention.wk.weight"], n_kv, n_kv * hd, dim)
        out[q + "self_attn.v_proj.weight"] = sd[p + "attention.wv.weight"]
        out[q + "self_attn.o_proj.weight"] = sd[p + "attention.wo.weight"]
        out[q + "mlp.gate_proj.weight"] = sd[p + "feed_forward.w1.weight"]
        out[q + "mlp.down_proj.weight"] = sd[p + "feed_forward.w2.weight"]
        out[q + "mlp.up_proj.weight"] = sd[p + "feed_forward.w3.weight"]
        out[q + "input_layernorm.weight"] = sd[p + "attention_norm.weight"]
        out[q + "post_attention_layernorm.weight"] = sd[p + "ffn_norm.weight"]
    inter = out["model.layers.0.mlp.gate_proj.weight"].shape[0]
    cfg = LlamaConfig(hidden_size=dim, intermediate_size=inter, num_attention_heads=n_heads,
                      num_key_value_heads=n_kv, num_hidden_layers=n_layers, rms_norm_eps=params["norm_eps"],
                      vocab_size=out["lm_head.weight"].shape[0],
                      max_position_embeddings=2048 if "n_kv_heads" not in params else 4096)
    with torch.device("meta"):
        model = LlamaForCausalLM(cfg)
    model.load_state_dict(out, assign=True, strict=True)
    model.config.torch_dtype = torch.float16
    model.save_pretrained(model_path, safe_serialization=safe_serialization)


def write_tokenizer(tokenizer_path, input_tokenizer_path):
    os.makedirs(tokenizer_path, exist_ok=True)
    try:
        from megatron2hf import load_sentencepiece_as_hf
        tok = load_sentencepiece_as_hf(str(input_tokenizer_path))
        assert len(tok) >= 10
        tok.save_pretrained(tokenizer_path)
    except Exception:  # noqa: BLE001 - without the HF tokenizer stack keep the sentencepiece model
        shutil.copy(input_tokenizer_path, os.path.join(tokenizer_path, "tokenizer.model"))


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--input_dir", required=True, help="directory with the Meta weights and tokenizer.model")
    p.add_argument("--model_size", type=int, choices=sorted(NUM_SHARDS), default=7)
    p.add_argument("--output_dir", required=True)
    p.add_argument("--unsafe", action="store_true", help="write .bin instead of safetensors")
    p.add_argument("--tokenizer_only", action="store_true")
    a = p.parse_args()
    if not a.tokenizer_only:
        write_model(a.output_dir, a.input_dir, a.model_size, safe_serialization=not a.unsafe)
    tok = os.path.join(a.input_dir, "tokenizer.model")
    if os.path.exists(tok):
        write_tokenizer(a.output_dir, tok)


if __name__ == "__main__":
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    main()
