"""HF Falcon / Meta-or-HF Llama -> single-rank Megatron 'release' checkpoint
(``<out>/release/mp_rank_00/model_optim_rng.pt`` + ``latest_checkpointed_iteration.txt`` =
``release``) carrying an ``args`` Namespace, ready for tools/checkpoint_util.py and
``finetune.py --load ... --use_checkpoint_args`` (reference: weights2megatron.py:16-261).

    python weights2megatron/weights2megatron.py llama2 --size 7 --model-path <hf dir> --out <dir>
"""
import os
import sys
from argparse import ArgumentParser, Namespace
from collections import OrderedDict
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
from epfl_megatron_b200.models.enums import PositionEmbeddingType  # noqa: E402
from epfl_megatron_b200.utils.ckpt_convert import fuse_qkv, permute_qkv, rope_hf_to_interleaved  # noqa: E402
from merge_llama import hf_state_to_meta, merge_llama  # noqa: E402

LLAMA_SIZES = {7: (32, 4096, 32, 32, 11008), 13: (40, 5120, 40, 40, 13824), 30: (60, 6656, 52, 52, 17920),
               34: (48, 8192, 64, 8, 22016), 65: (80, 8192, 64, 64, 22016), 70: (80, 8192, 64, 8, 28672)}
FALCON_SIZES = {7: (32, 4544, 71, 1), 40: (60, 8192, 128, 8)}


def llama_to_megatron(weights, n_layer, hidden, n_heads, n_kv, source="hf"):
    """Meta-named Llama weights -> Megatron language_model dict (modern key spelling)."""
    hn = hidden // n_heads
    enc = OrderedDict()
    for i in range(n_layer):
        p = f"layers.{i}"
        wq, wk, wv = (weights[f"{p}.attention.w{x}.weight"] for x in "qkv")
        if source == "hf":  # HF stores q/k in half-rotation order; Megatron wants interleaved pairs
            wq = rope_hf_to_interleaved(wq, n_heads, hn)
            wk = rope_hf_to_interleaved(wk, n_kv, hn)
        enc[f"{p}.self_attention.query_key_value.weight"] = fuse_qkv(wq, wk, wv, n_heads, n_kv)
        enc[f"{p}.self_attention.dense.weight"] = weights[f"{p}.attention.wo.weight"]
        # GLU weight = [up (w3); gate (w1)]: the activation computes x1 * silu(x2)
        enc[f"{p}.mlp.dense_h_to_4h.weight"] = torch.cat(
            [weights[f"{p}.feed_forward.w3.weight"], weights[f"{p}.feed_forward.w1.weight"]], dim=0)
        enc[f"{p}.mlp.dense_4h_to_h.weight"] = weights[f"{p}.feed_forward.w2.weight"]
        enc[f"{p}.input_layernorm.weight"] = weights[f"{p}.attention_norm.weight"]
        enc[f"{p}.post_attention_layernorm.weight"] = weights[f"{p}.ffn_norm.weight"]
    enc["final_layernorm.weight"] = weights["norm.weight"]
    return {"embedding": {"word_embeddings": {"weight": weights["tok_embeddings.weight"]}},
            "encoder": enc, "lm_head": weights["output.weight"]}


def falcon_to_megatron(sd, n_layer, hidden, n_heads, n_kv, parallel_layernorm):
    """HF Falcon state dict -> Megatron language_model dict. HF's fused QKV already has the
    per-KV-group layout; only the RoPE row order of q and k changes."""
    enc = OrderedDict()
    for i in range(n_layer):
        hp, p = f"transformer.h.{i}", f"layers.{i}"
        enc[f"{p}.self_attention.query_key_value.weight"] = permute_qkv(
            sd[f"{hp}.self_attention.query_key_value.weight"], hidden, n_heads, n_kv)
        enc[f"{p}.self_attention.dense.weight"] = sd[f"{hp}.self_attention.dense.weight"]
        enc[f"{p}.mlp.dense_h_to_4h.weight"] = sd[f"{hp}.mlp.dense_h_to_4h.weight"]
        enc[f"{p}.mlp.dense_4h_to_h.weight"] = sd[f"{hp}.mlp.dense_4h_to_h.weight"]
        if parallel_layernorm:
            for x in ("weight", "bias"):
                enc[f"{p}.input_layernorm.{x}"] = sd[f"{hp}.ln_attn.{x}"]
                enc[f"{p}.mlp_layernorm.{x}"] = sd[f"{hp}.ln_mlp.{x}"]
        else:
            for x in ("weight", "bias"):
                enc[f"{p}.input_layernorm.{x}"] = sd[f"{hp}.input_layernorm.{x}"]
    enc["final_layernorm.weight"] = sd["transformer.ln_f.weight"]
    enc["final_layernorm.bias"] = sd["transformer.ln_f.bias"]
    return {"embedding": {"word_embeddings": {"weight": sd["transformer.word_embeddings.weight"]}},
            "encoder": enc}


def make_args(model, lm, n_layer, hidden, n_heads, n_kv, ffn, seq, extra):
    vocab = lm["embedding"]["word_embeddings"]["weight"].shape[0]
    a = dict(num_layers=n_layer, hidden_size=hidden, num_attention_heads=n_heads, num_attention_heads_kv=n_kv,
             ffn_hidden_size=ffn, kv_channels=hidden // n_heads, parallel_attn=False, parallel_layernorm=False,
             make_vocab_size_divisible_by=1, padded_vocab_size=vocab, glu_activation=None, use_rms_norm=False,
             tie_embed_logits=True, tokenizer_type="FalconTokenizer", max_position_embeddings=seq,
             seq_length=seq, layernorm_epsilon=1e-5, rope_scaling_factor=1.0,
             position_embedding_type=PositionEmbeddingType.rotary, use_bias=False, use_post_ln=False,
             tensor_model_parallel_size=1, pipeline_model_parallel_size=1, iteration="release",
             bias_gelu_fusion=False, bias_dropout_fusion=False, hidden_dropout=0.0, attention_dropout=0.0,
             consumed_train_samples=0, consumed_valid_samples=0, params_dtype=torch.float32,
             use_distributed_optimizer=False, model_name=model)
    a.update(extra)
    return Namespace(**a)


def save_release(lm, args, out_dir, dtype):
    def cast(x):
        if isinstance(x, dict):
            return type(x)((k, cast(v)) for k, v in x.items())
        return x.to(dtype).clone() if torch.is_floating_point(x) else x
    # reference checkpoints pickle megatron.model.enums; make ours loadable by the reference's tools too
    try:
        import megatron.model.enums  # noqa: F401  (facade)
    except ImportError:
        pass
    out = Path(out_dir)
    (out / "release" / "mp_rank_00").mkdir(parents=True, exist_ok=True)
    torch.save({"iteration": "release", "model": {"language_model": cast(lm)}, "checkpoint_version": 3.0,
                "args": args}, out / "release" / "mp_rank_00" / "model_optim_rng.pt")
    (out / "latest_checkpointed_iteration.txt").write_text("release")
    print("saved release checkpoint to", out)


def convert(model, size, out, model_path=None, cache_dir=None, dtype=torch.bfloat16, state_dict=None,
            config=None):
    """``state_dict``/``config`` allow in-memory conversion (tests, no files)."""
    if model == "falcon":
        if state_dict is None:
            from transformers import AutoModelForCausalLM
            hf = AutoModelForCausalLM.from_pretrained(model_path or f"tiiuae/falcon-{size}b", cache_dir=cache_dir)
            state_dict, config = hf.state_dict(), hf.config
        n_layer, hidden, n_heads = config.num_hidden_layers, config.hidden_size, config.num_attention_heads
        new_arch = getattr(config, "new_decoder_architecture", False)
        n_kv = config.num_kv_heads if new_arch else (1 if getattr(config, "multi_query", True) else n_heads)
        lm = falcon_to_megatron(state_dict, n_layer, hidden, n_heads, n_kv, new_arch)
        args = make_args("falcon", lm, n_layer, hidden, n_heads, n_kv, 4 * hidden, 2048,
                         dict(parallel_attn=True, parallel_layernorm=bool(new_arch), tie_embed_logits=True,
                              tokenizer_type="FalconTokenizer"))
    else:
        version = 1 if model == "llama" else 2
        if state_dict is None:
            weights, source = merge_llama(size, version, cache_dir, model_path)
            n_layer, hidden, n_heads, n_kv, ffn = LLAMA_SIZES[size]
            # the size table covers the published checkpoints; when the source directory describes
            # itself (HF config.json, Meta params.json) its numbers win, so any Llama-shaped model converts
            src = Path(model_path or cache_dir or ".")
            if source == "hf" and (src / "config.json").exists():
                from transformers import AutoConfig
                config = AutoConfig.from_pretrained(src)
                n_layer, hidden, n_heads = config.num_hidden_layers, config.hidden_size, config.num_attention_heads
                n_kv, ffn = config.num_key_value_heads, config.intermediate_size
            elif source == "meta" and (src / "params.json").exists():
                import json
                params = json.loads((src / "params.json").read_text())
                n_layer, hidden, n_heads = params["n_layers"], params["dim"], params["n_heads"]
                n_kv = params.get("n_kv_heads", n_heads)
                ffn = weights["layers.0.feed_forward.w1.weight"].shape[0]
        else:
            weights, source = hf_state_to_meta(state_dict), "hf"
            n_layer, hidden, n_heads = config.num_hidden_layers, config.hidden_size, config.num_attention_heads
            n_kv, ffn = config.num_key_value_heads, config.intermediate_size
        lm = llama_to_megatron(weights, n_layer, hidden, n_heads, n_kv, source)
        seq = 2048 if version == 1 else (16384 if model == "codellama" else 4096)
        if config is not None:
            seq = config.max_position_embeddings
        args = make_args(model, lm, n_layer, hidden, n_heads, n_kv, ffn, seq,
                         dict(glu_activation="swiglu", use_rms_norm=True, tie_embed_logits=False,
                              tokenizer_type="SentencePieceTokenizer",
                              layernorm_epsilon=1e-6 if version == 1 else 1e-5))
    save_release(lm, args, out, dtype)
    return lm, args


def main():
    p = ArgumentParser(description="Convert Huggingface llama or falcon weights to megatron-compatible weights")
    p.add_argument("model", choices={"falcon", "llama", "llama2", "codellama"})
    p.add_argument("--size", default=7, choices={7, 13, 30, 34, 40, 65, 70}, type=int)
    p.add_argument("--out", type=Path, required=True)
    p.add_argument("--cache-dir", "--cache_dir", type=Path)
    p.add_argument("--model-path", "--model_path", type=str)
    p.add_argument("--dtype", default="bf16", choices=["bf16", "fp16", "fp32"])
    a = p.parse_args()
    dt = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[a.dtype]
    convert(a.model, a.size, a.out, a.model_path, a.cache_dir, dt)


if __name__ == "__main__":
    main()
