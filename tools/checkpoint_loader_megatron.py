"""Loader plug-in of tools/checkpoint_util.py (tools/checkpoint_loader_megatron.py:21-346
surface: ``add_arguments(parser)`` + ``load_checkpoint(queue, args)``).

Protocol on the queue (named messages, as in the reference's :24-86 description): first a
metadata namespace, then one message per piece of the model -- "embeddings", "transformer
layer <n>", "final layernorm", "lm_head" (untied models) -- each a dict of full, un-sharded
tensors, and finally the string "done". A failure sends "exit". The tensors come from merging
the checkpoint FILES (utils/ckpt_convert), so no model is built and no GPU is needed.
"""
import os
import sys
import types

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))


def add_arguments(parser):
    group = parser.add_argument_group(title="Megatron loader")
    group.add_argument("--true_vocab_size", type=int, default=None,
                       help="original size of vocab, if specified will trim padding from embedding table.")
    group.add_argument("--vocab_file", type=str, default=None,
                       help="Path to the vocab file. If specified will use this to get vocab size and trim padding")
    group.add_argument("--megatron_path", type=str, default=None, help="Base directory of the framework")


def _load_checkpoint(queue, args):
    from tools.checkpoint_util import load_full
    from epfl_megatron_b200.utils.ckpt_convert import layer_index
    full, src_args, iteration = load_full(args.load_dir)
    md = types.SimpleNamespace()
    md.model_type = args.model_type
    for k in ("num_layers", "hidden_size", "seq_length", "num_attention_heads", "num_attention_heads_kv",
              "max_position_embeddings", "tokenizer_type", "params_dtype", "ffn_hidden_size", "glu_activation",
              "tie_embed_logits", "parallel_attn", "parallel_layernorm", "use_rms_norm", "use_bias",
              "position_embedding_type", "padded_vocab_size", "make_vocab_size_divisible_by"):
        setattr(md, k, getattr(src_args, k, None))
    md.iteration = iteration
    md.previous_tensor_parallel_size = getattr(src_args, "tensor_model_parallel_size", 1)
    md.previous_pipeline_parallel_size = getattr(src_args, "pipeline_model_parallel_size", 1)
    md.true_vocab_size = args.true_vocab_size
    md.checkpoint_args = src_args
    md.consumed_train_samples = getattr(src_args, "consumed_train_samples", 0)
    md.consumed_valid_samples = getattr(src_args, "consumed_valid_samples", 0)
    queue.put(md)

    def send(name, msg):
        print(f"sending {name}")
        msg["name"] = name
        queue.put(msg)

    emb = full["embedding"]
    msg = {"word embeddings": emb["word_embeddings"]["weight"]}
    if "position_embeddings" in emb:
        msg["position embeddings"] = emb["position_embeddings"]["weight"]
    if "tokentype_embeddings" in emb:
        msg["tokentype embeddings"] = emb["tokentype_embeddings"]["weight"]
    send("embeddings", msg)
    enc = full["encoder"]
    n_layers = 1 + max((layer_index(k)[0] for k in enc if layer_index(k)[0] is not None), default=-1)
    for i in range(n_layers):
        pre = f"layers.{i}."
        send(f"transformer layer {i}", {k[len(pre):]: v for k, v in enc.items() if k.startswith(pre)})
    send("final layernorm", {k[len("final_layernorm."):]: v for k, v in enc.items()
                             if k.startswith("final_layernorm.")})
    if "lm_head" in full:
        send("lm_head", {"lm_head": full["lm_head"]})
    if "pooler" in full:
        send("pooler", dict(full["pooler"]))
    for name, sub_d in full.get("heads", {}).items():      # BERT: "lm_head" (module), "binary_head"
        send(f"head {name}", dict(sub_d))
    queue.put("done")


def load_checkpoint(queue, args):
    try:
        _load_checkpoint(queue, args)
    except Exception:
        queue.put("exit")
        raise
