"""Re-shard a Megatron checkpoint to a different tensor / pipeline parallel layout.

    python tools/checkpoint_util.py --model_type llama2 --load_dir <in> --save_dir <out> \
        --target_tensor_parallel_size 4 --target_pipeline_parallel_size 2 [--true_vocab_size 32000]

Same CLI as the reference (tools/checkpoint_util.py:106-148 with its loader/saver plug-ins
checkpoint_loader_megatron.py / checkpoint_saver_megatron.py). The reference builds tp x pp
CPU models with faked ranks in one process and streams tensors to a saver process over a
queue; here the checkpoint FILES are merged and re-split directly with the layout rules in
``epfl_megatron_b200.utils.ckpt_convert`` (fused-QKV per KV group on dim 0, GLU
[up_shard; gate_shard], row-parallel on dim 1, vocab re-padding), which needs no model
construction, no process group and no GPU. Optimizer state is not carried over (as in the
reference).
"""
import argparse
import copy
import os
import sys
from collections import OrderedDict
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from epfl_megatron_b200.utils.ckpt_convert import (layer_index, merge_tp, normalize_language_model,  # noqa: E402
                                                   pad_vocab, padded_vocab_size, split_tp)


def _tracker(load_dir):
    with open(os.path.join(load_dir, "latest_checkpointed_iteration.txt")) as f:
        it = f.read().strip()
    if it != "release":
        it = int(it)            # the training loop compares the stored iteration with integers
    return it, ("release" if it == "release" else f"iter_{it:07d}")


def _rank_dir(tp_rank, pp_rank, pp):
    return f"mp_rank_{tp_rank:02d}" if pp == 1 else f"mp_rank_{tp_rank:02d}_{pp_rank:03d}"


def _load_file(d):
    for name in ("model_optim_rng.pt", "model_rng.pt"):
        p = os.path.join(d, name)
        if os.path.isfile(p):
            try:
                import megatron  # noqa: F401  (facade for unpickling)
            except ImportError:
                pass
            return torch.load(p, map_location="cpu", weights_only=False)
    raise FileNotFoundError(f"no model file under {d}")


def load_full(load_dir):
    """-> (full language_model dict with global layer numbering, source args, iteration)."""
    it, sub = _tracker(load_dir)
    root = os.path.join(load_dir, sub)
    first = None
    for cand in ("mp_rank_00", "mp_rank_00_000"):
        if os.path.isdir(os.path.join(root, cand)):
            first = _load_file(os.path.join(root, cand))
            break
    assert first is not None, f"no mp_rank_00[_000] under {root}"
    a = first["args"]
    tp = getattr(a, "tensor_model_parallel_size", 1)
    pp = getattr(a, "pipeline_model_parallel_size", 1)
    glu = getattr(a, "glu_activation", None)
    L = a.num_layers
    full = {"embedding": OrderedDict(), "encoder": OrderedDict()}
    for p in range(pp):
        shards = [_load_file(os.path.join(root, _rank_dir(t, p, pp))) for t in range(tp)]
        chunks = [k for k in shards[0] if k == "model" or (k.startswith("model") and k[5:].isdigit())]
        chunks.sort(key=lambda k: int(k[5:]) if len(k) > 5 else 0)
        vpp = len(chunks)
        for c, ck in enumerate(chunks):
            lms = [normalize_language_model(s[ck]["language_model"]) for s in shards]
            if "embedding" in lms[0] and not full["embedding"]:
                for name, sub_d in lms[0]["embedding"].items():
                    full["embedding"][name] = OrderedDict()
                    for leaf in sub_d:
                        parts = [lm["embedding"][name][leaf] for lm in lms]
                        full["embedding"][name][leaf] = torch.cat(parts, dim=0) if name == "word_embeddings" \
                            else parts[0].clone()
            enc = lms[0].get("encoder", {})
            n_local = len({layer_index(k)[0] for k in enc if layer_index(k)[0] is not None})
            for k in enc:
                li, rest = layer_index(k)
                parts = [lm["encoder"][k] for lm in lms]
                if li is None:
                    full["encoder"][k] = merge_tp(k, parts, glu)
                else:
                    g = c * (L // vpp) + p * n_local + li if vpp > 1 else p * n_local + li
                    full["encoder"][f"layers.{g}.{rest}"] = merge_tp(rest, parts, glu)
            if "lm_head" in lms[0]:
                full["lm_head"] = torch.cat([lm["lm_head"] for lm in lms], dim=0)
            if "pooler" in lms[0]:
                full["pooler"] = lms[0]["pooler"]
            # task heads next to the language model (BERT: lm_head, binary_head; classification /
            # multiple-choice heads): replicated over TP except the vocabulary-parallel lm_head bias
            for name, sub_d in shards[0][ck].items():
                if name in ("language_model", "word_embeddings_for_head") or not isinstance(sub_d, dict):
                    continue
                full.setdefault("heads", OrderedDict())[name] = OrderedDict(
                    (leaf, torch.cat([s[ck][name][leaf] for s in shards], dim=0)
                     if (name, leaf) == ("lm_head", "bias") else val.clone())
                    for leaf, val in sub_d.items() if torch.is_tensor(val))
    return full, a, it


def save_sharded(full, src_args, iteration, save_dir, tp, pp, true_vocab_size=None):
    a = copy.deepcopy(src_args)
    glu = getattr(a, "glu_activation", None)
    L = a.num_layers
    assert L % pp == 0, "num_layers must be divisible by the target pipeline size"
    assert a.num_attention_heads % tp == 0 and a.num_attention_heads_kv % tp == 0, \
        "attention heads (and KV heads) must be divisible by the target tensor parallel size"
    div = getattr(a, "make_vocab_size_divisible_by", 128)
    emb = full["embedding"]["word_embeddings"]["weight"]
    true_v = true_vocab_size or emb.shape[0]
    new_v = padded_vocab_size(true_v, div, tp)
    emb = pad_vocab(emb[:true_v], new_v)
    head = pad_vocab(full["lm_head"][:true_v], new_v) if "lm_head" in full else None
    a.tensor_model_parallel_size, a.pipeline_model_parallel_size = tp, pp
    a.padded_vocab_size = new_v
    sub = "release" if iteration == "release" else f"iter_{int(iteration):07d}"
    per = L // pp
    tied = getattr(a, "tie_embed_logits", head is None)
    for p in range(pp):
        enc_full = OrderedDict()
        for k, v in full["encoder"].items():
            li, rest = layer_index(k)
            if li is None:
                if p == pp - 1:
                    enc_full[k] = v
            elif p * per <= li < (p + 1) * per:
                enc_full[f"layers.{li - p * per}.{rest}"] = v
        enc_split = {k: split_tp(layer_index(k)[1] if layer_index(k)[0] is not None else k, v, tp, glu)
                     for k, v in enc_full.items()}
        emb_split = emb.chunk(tp, dim=0)
        head_split = head.chunk(tp, dim=0) if head is not None else None
        for t in range(tp):
            lm = OrderedDict()
            if p == 0:
                e = OrderedDict()
                for name, sub_d in full["embedding"].items():
                    e[name] = OrderedDict((leaf, emb_split[t].clone() if name == "word_embeddings" else val.clone())
                                          for leaf, val in sub_d.items())
                lm["embedding"] = e
            lm["encoder"] = OrderedDict((k, v[t]) for k, v in enc_split.items())
            if p == pp - 1 and head_split is not None:
                lm["lm_head"] = head_split[t].clone()
            if p == pp - 1 and "pooler" in full:
                lm["pooler"] = full["pooler"]
            model = {"language_model": lm}
            if p == pp - 1:
                for name, sub_d in full.get("heads", {}).items():
                    model[name] = OrderedDict()
                    for leaf, val in sub_d.items():
                        if (name, leaf) == ("lm_head", "bias"):
                            val = torch.cat([val[:true_v], val.new_zeros(max(new_v - true_v, 0))])[:new_v]
                            val = val.chunk(tp, dim=0)[t]
                        model[name][leaf] = val.clone()
            if tied and pp > 1 and p == pp - 1:
                model["word_embeddings_for_head"] = OrderedDict(weight=emb_split[t].clone())
            out = os.path.join(save_dir, sub, _rank_dir(t, p, pp))
            os.makedirs(out, exist_ok=True)
            torch.save({"args": a, "checkpoint_version": 3.0, "iteration": iteration, "model": model},
                       os.path.join(out, "model_optim_rng.pt"))
    with open(os.path.join(save_dir, "latest_checkpointed_iteration.txt"), "w") as f:
        f.write(str(iteration))


def main():
    p = argparse.ArgumentParser(description="Megatron checkpoint utility: change tp / pp of a checkpoint",
                                allow_abbrev=False)
    p.add_argument("--model_type", type=str, default="GPT",
                   choices=["GPT", "BERT", "falcon", "llama", "llama2", "codellama"])
    p.add_argument("--loader", type=str, default="megatron")
    p.add_argument("--saver", type=str, default="megatron")
    p.add_argument("--load_dir", type=str, required=True)
    p.add_argument("--save_dir", type=str, required=True)
    p.add_argument("--max_queue_size", type=int, default=50)
    p.add_argument("--no_checking", action="store_false", dest="checking")
    p.add_argument("--target_tensor_parallel_size", type=int, default=1)
    p.add_argument("--target_pipeline_parallel_size", type=int, default=1)
    p.add_argument("--true_vocab_size", type=int, default=None)
    p.add_argument("--vocab_file", type=str, default=None)
    p.add_argument("--megatron_path", type=str, default=None)
    p.add_argument("--bf16", action="store_true")
    p.add_argument("--fp16", action="store_true")
    a = p.parse_args()
    full, src_args, it = load_full(a.load_dir)
    if a.bf16 or a.fp16:
        dt = torch.bfloat16 if a.bf16 else torch.float16

        def cast(x):
            if isinstance(x, dict):
                return type(x)((k, cast(v)) for k, v in x.items())
            return x.to(dt) if torch.is_floating_point(x) else x
        full = cast(full)
    save_sharded(full, src_args, it, a.save_dir, a.target_tensor_parallel_size, a.target_pipeline_parallel_size,
                 a.true_vocab_size)
    print(f"saved tp={a.target_tensor_parallel_size} pp={a.target_pipeline_parallel_size} checkpoint to {a.save_dir}")


if __name__ == "__main__":
    main()
