"""Saver plug-in of tools/checkpoint_util.py (tools/checkpoint_saver_megatron.py:19-402
surface: ``add_arguments(parser)`` + ``save_checkpoint(queue, args)``). Consumes the messages
of checkpoint_loader_megatron.py, reassembles the full model and writes it at the target
tensor / pipeline parallel sizes (vocab re-padding, per-KV-group QKV split, per-rank
[up_shard; gate_shard] GLU layout) with the regular checkpoint naming."""
import os
import sys
from collections import OrderedDict

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))


def add_arguments(parser):
    group = parser.add_argument_group(title="Megatron saver")
    group.add_argument("--megatron_path", type=str, default=None, help="Base directory of the framework")
    group.add_argument("--target_tensor_parallel_size", type=int,
                       help="Target tensor model parallel size, defaults to the tensor parallel size "
                            "in the input checkpoint if provided by the loader, otherwise to 1")
    group.add_argument("--target_pipeline_parallel_size", type=int,
                       help="Target pipeline model parallel size, default to the pipeline parallel size "
                            "in the input checkpoint if provided by the loader, otherwise to 1")


def save_checkpoint(queue, args):
    from tools.checkpoint_util import save_sharded

    def queue_get(name=None):
        val = queue.get()
        if val == "exit":
            print("Loader exited, exiting saver")
            sys.exit(1)
        if name is not None and args.checking and val["name"] != name:
            print(f'Unexpected message. Expecting "{name}" but got "{val["name"]}". Exiting saver.')
            sys.exit(1)
        if name is not None:
            print(f"received {name}")
        return val

    def check_message(msg):
        if not args.checking:
            return
        msg.pop("name", None)

    md = queue_get()
    tp = args.target_tensor_parallel_size or getattr(md, "previous_tensor_parallel_size", 1) or 1
    pp = args.target_pipeline_parallel_size or getattr(md, "previous_pipeline_parallel_size", 1) or 1
    full = OrderedDict()
    emb = queue_get("embeddings")
    full["embedding"] = OrderedDict(word_embeddings=OrderedDict(weight=emb["word embeddings"]))
    if "position embeddings" in emb:
        full["embedding"]["position_embeddings"] = OrderedDict(weight=emb["position embeddings"])
    if "tokentype embeddings" in emb:
        full["embedding"]["tokentype_embeddings"] = OrderedDict(weight=emb["tokentype embeddings"])
    enc = OrderedDict()
    for i in range(md.num_layers):
        msg = queue_get(f"transformer layer {i}")
        check_message(msg)
        for k, v in msg.items():
            if k != "name":
                enc[f"layers.{i}.{k}"] = v
    msg = queue_get("final layernorm")
    for k, v in msg.items():
        if k != "name":
            enc[f"final_layernorm.{k}"] = v
    full["encoder"] = enc
    msg = queue_get()
    while msg != "done":
        name = msg.pop("name", None)
        if name == "lm_head":
            full["lm_head"] = msg["lm_head"]
        elif name == "pooler":
            full["pooler"] = OrderedDict(msg)
        elif name is not None and name.startswith("head "):
            full.setdefault("heads", OrderedDict())[name[len("head "):]] = OrderedDict(msg)
        else:
            print(f"ERROR: unexpected message {name!r}; was expecting to be done")
            break
        msg = queue_get()
    true_vocab = getattr(args, "true_vocab_size", None) or getattr(md, "true_vocab_size", None)
    save_sharded(full, md.checkpoint_args, md.iteration, args.save_dir, tp, pp, true_vocab)
    print(f"Done! saved tp={tp} pp={pp} checkpoint to {args.save_dir}")
