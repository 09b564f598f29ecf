"""Unsharded Megatron checkpoint (tp = pp = 1) -> Hugging Face ``LlamaForCausalLM`` /
``FalconForCausalLM`` directory, optionally with the tokenizer
(reference: weights2megatron/megatron2hf.py:60-431).

    python weights2megatron/megatron2hf.py --model llama2 --input_dir <megatron ckpt> --output_dir <hf dir>
"""
import json
import os
import sys
from argparse import ArgumentParser
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from epfl_megatron_b200.utils.ckpt_convert import (normalize_language_model, permute_qkv,  # noqa: E402
                                                   rope_interleaved_to_hf, split_qkv)


def load_megatron(input_dir):
    input_dir = Path(input_dir)
    with open(input_dir / "latest_checkpointed_iteration.txt") as f:
        it = f.read().strip()
    sub = "release" if it == "release" else f"iter_{int(it):07d}"
    base = input_dir / sub / "mp_rank_00"
    path = base / "model_optim_rng.pt"
    if not path.exists():
        path = base / "model_rng.pt"
    try:
        import megatron  # noqa: F401  facade: lets torch.load unpickle megatron.* classes
    except ImportError:
        pass
    sd = torch.load(path, map_location="cpu", weights_only=False)
    args = sd["args"]
    assert getattr(args, "tensor_model_parallel_size", 1) == 1 and \
        getattr(args, "pipeline_model_parallel_size", 1) == 1, \
        "unshard the checkpoint first (tools/checkpoint_util.py --target_tensor_parallel_size 1 ...)"
    return normalize_language_model(sd["model"]["language_model"]), args


def convert_wqkv(wqkv, n_heads, n_kv):
    """Grouped fused QKV (interleaved RoPE rows) -> HF q, k, v (half-rotation rows)."""
    hn = wqkv.shape[0] // (n_heads + 2 * n_kv)
    wq, wk, wv = split_qkv(wqkv, n_heads, n_kv)
    return rope_interleaved_to_hf(wq, n_heads, hn), rope_interleaved_to_hf(wk, n_kv, hn), wv


def convert_ffn(w):
    """[up (w3); gate (w1)] -> (gate, up)."""
    up, gate = w.chunk(2, dim=0)
    return gate, up


def llama_state_dict(lm, args, true_vocab_size=None):
    enc = lm["encoder"]
    out = {}
    emb = lm["embedding"]["word_embeddings"]["weight"]
    head = lm["lm_head"]
    if true_vocab_size:
        emb, head = emb[:true_vocab_size], head[:true_vocab_size]
    out["model.embed_tokens.weight"] = emb
    out["lm_head.weight"] = head
    out["model.norm.weight"] = enc["final_layernorm.weight"]
    for i in range(args.num_layers):
        p, h = f"layers.{i}", f"model.layers.{i}"
        q, k, v = convert_wqkv(enc[f"{p}.self_attention.query_key_value.weight"], args.num_attention_heads,
                               args.num_attention_heads_kv)
        out[f"{h}.self_attn.q_proj.weight"], out[f"{h}.self_attn.k_proj.weight"] = q, k
        out[f"{h}.self_attn.v_proj.weight"] = v
        out[f"{h}.self_attn.o_proj.weight"] = enc[f"{p}.self_attention.dense.weight"]
        gate, up = convert_ffn(enc[f"{p}.mlp.dense_h_to_4h.weight"])
        out[f"{h}.mlp.gate_proj.weight"], out[f"{h}.mlp.up_proj.weight"] = gate, up
        out[f"{h}.mlp.down_proj.weight"] = enc[f"{p}.mlp.dense_4h_to_h.weight"]
        out[f"{h}.input_layernorm.weight"] = enc[f"{p}.input_layernorm.weight"]
        out[f"{h}.post_attention_layernorm.weight"] = enc[f"{p}.post_attention_layernorm.weight"]
    return {k: v.clone().contiguous() for k, v in out.items()}


def llama_config(args, vocab_size, rope_scaling=None):
    from transformers import LlamaConfig
    cfg = LlamaConfig(vocab_size=vocab_size, hidden_size=args.hidden_size, intermediate_size=args.ffn_hidden_size,
                      num_hidden_layers=args.num_layers, num_attention_heads=args.num_attention_heads,
                      num_key_value_heads=args.num_attention_heads_kv,
                      max_position_embeddings=args.max_position_embeddings, rms_norm_eps=args.layernorm_epsilon,
                      tie_word_embeddings=False, attention_bias=False)
    sf = getattr(args, "rope_scaling_factor", 1.0)
    if sf and sf != 1.0:
        cfg.rope_scaling = {"type": "linear", "factor": sf, "rope_type": "linear"}
    return cfg


def write_llama_model(model_path, input_dir, true_vocab_size=None, dtype=torch.bfloat16, safe=True):
    from transformers import LlamaForCausalLM
    lm, args = load_megatron(input_dir)
    sd = llama_state_dict(lm, args, true_vocab_size)
    cfg = llama_config(args, sd["model.embed_tokens.weight"].shape[0])
    model = LlamaForCausalLM(cfg).to(dtype)
    missing, unexpected = model.load_state_dict({k: v.to(dtype) for k, v in sd.items()}, strict=False)
    assert not [m for m in missing if "rotary" not in m], missing
    assert not unexpected, unexpected
    model.save_pretrained(model_path, safe_serialization=safe)
    return model


def falcon_state_dict(lm, args):
    enc = lm["encoder"]
    out = {"transformer.word_embeddings.weight": lm["embedding"]["word_embeddings"]["weight"],
           "lm_head.weight": lm["embedding"]["word_embeddings"]["weight"],
           "transformer.ln_f.weight": enc["final_layernorm.weight"],
           "transformer.ln_f.bias": enc["final_layernorm.bias"]}
    for i in range(args.num_layers):
        p, h = f"layers.{i}", f"transformer.h.{i}"
        out[f"{h}.self_attention.query_key_value.weight"] = permute_qkv(
            enc[f"{p}.self_attention.query_key_value.weight"], args.hidden_size, args.num_attention_heads,
            args.num_attention_heads_kv, revert=True)
        out[f"{h}.self_attention.dense.weight"] = enc[f"{p}.self_attention.dense.weight"]
        out[f"{h}.mlp.dense_h_to_4h.weight"] = enc[f"{p}.mlp.dense_h_to_4h.weight"]
        out[f"{h}.mlp.dense_4h_to_h.weight"] = enc[f"{p}.mlp.dense_4h_to_h.weight"]
        if getattr(args, "parallel_layernorm", False):
            for x in ("weight", "bias"):
                out[f"{h}.ln_attn.{x}"] = enc[f"{p}.input_layernorm.{x}"]
                out[f"{h}.ln_mlp.{x}"] = enc[f"{p}.mlp_layernorm.{x}"]
        else:
            for x in ("weight", "bias"):
                out[f"{h}.input_layernorm.{x}"] = enc[f"{p}.input_layernorm.{x}"]
    return {k: v.clone().contiguous() for k, v in out.items()}


def write_falcon_model(model_path, input_dir, dtype=torch.bfloat16, safe=True):
    from transformers import FalconConfig, FalconForCausalLM
    lm, args = load_megatron(input_dir)
    sd = falcon_state_dict(lm, args)
    new_arch = bool(getattr(args, "parallel_layernorm", False))
    cfg = FalconConfig(vocab_size=sd["transformer.word_embeddings.weight"].shape[0], hidden_size=args.hidden_size,
                       num_hidden_layers=args.num_layers, num_attention_heads=args.num_attention_heads,
                       num_kv_heads=args.num_attention_heads_kv, new_decoder_architecture=new_arch,
                       multi_query=args.num_attention_heads_kv == 1, parallel_attn=True, bias=False,
                       alibi=False, layer_norm_epsilon=args.layernorm_epsilon)
    model = FalconForCausalLM(cfg).to(dtype)
    model.load_state_dict({k: v.to(dtype) for k, v in sd.items()}, strict=False)
    model.save_pretrained(model_path, safe_serialization=safe)
    return model


def load_sentencepiece_as_hf(vocab_file):
    """HF Llama tokenizer over a SentencePiece model file. transformers 5 reads the model only through
    ``from_pretrained`` on a directory holding ``tokenizer.model`` (its constructor's first arguments became
    vocab / merges); transformers 4 took ``vocab_file=`` directly."""
    import shutil
    import tempfile
    from transformers import LlamaTokenizer
    with tempfile.TemporaryDirectory() as d:
        shutil.copy(vocab_file, os.path.join(d, "tokenizer.model"))
        try:
            tok = LlamaTokenizer.from_pretrained(d)
        except Exception:  # noqa: BLE001
            tok = None
    if tok is None or len(tok) < 10:
        tok = LlamaTokenizer(vocab_file=vocab_file)
    return tok


def write_tokenizer(args, model_path, vocab_file=None, override_special_tokens=None):
    """Save the tokenizer next to the model; Megatron's extra special tokens are added and the
    ids are checked against the Megatron tokenizer."""
    tt = getattr(args, "tokenizer_type", None)
    if tt == "FalconTokenizer":
        from transformers import AutoTokenizer
        tok = AutoTokenizer.from_pretrained(vocab_file or "tiiuae/falcon-40b")
    elif tt == "SentencePieceTokenizer" and vocab_file:
        tok = load_sentencepiece_as_hf(vocab_file)
        from epfl_megatron_b200.tokenizer.tokenizer import _SentencePieceTokenizer
        mt = _SentencePieceTokenizer(vocab_file, vocab_extra_ids=getattr(args, "vocab_extra_ids", 0),
                                     vocab_extra_ids_list=getattr(args, "vocab_extra_ids_list", None),
                                     new_tokens=getattr(args, "new_tokens", True))
        extra = [t for t, i in sorted(mt._special_tokens.items(), key=lambda kv: kv[1]) if i >= len(tok)]
        if extra:
            tok.add_special_tokens({"additional_special_tokens": extra})
        for t, i in mt._special_tokens.items():
            assert tok.convert_tokens_to_ids(t) == i, f"token id mismatch for {t}"
    else:
        print("no tokenizer written (unknown tokenizer type or no --vocab_file)")
        return None
    if override_special_tokens:
        tok.add_special_tokens(dict(kv.split("=") for kv in override_special_tokens))
    tok.save_pretrained(model_path)
    return tok


def main():
    p = ArgumentParser()
    p.add_argument("--input_dir", required=True)
    p.add_argument("--output_dir", required=True)
    p.add_argument("--model", choices={"falcon", "llama", "llama2", "codellama"}, default="llama2")
    p.add_argument("--vocab_file", type=str)
    p.add_argument("--vocab_extra_ids_list", type=str)
    p.add_argument("--true_vocab_size", type=int)
    p.add_argument("--no_new_tokens", action="store_false", dest="new_tokens")
    p.add_argument("--override_special_tokens", nargs="*", default=[])
    p.add_argument("--dtype", default="bf16", choices=["bf16", "fp16", "fp32"])
    a = p.parse_args()
    dt = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[a.dtype]
    if a.model == "falcon":
        write_falcon_model(a.output_dir, a.input_dir, dt)
    else:
        write_llama_model(a.output_dir, a.input_dir, a.true_vocab_size, dt)
    _, margs = load_megatron(a.input_dir)
    try:
        write_tokenizer(margs, a.output_dir, a.vocab_file, a.override_special_tokens)
    except OSError as e:
        # e.g. the Falcon tokenizer lives on the hub and this machine is offline: the weights are
        # exported, the tokenizer can be added later (--vocab_file <local tokenizer directory>)
        print(f"WARNING: model written, tokenizer NOT written ({type(e).__name__}: {str(e).splitlines()[0]})")


if __name__ == "__main__":
    main()
