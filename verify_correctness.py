"""Logit / loss parity between a checkpoint run through this framework and an independent
baseline (verify_correctness.py:50-194 surface): ``--load`` is a Megatron checkpoint directory
(or a Hugging Face one), ``--huggingface_cache`` the baseline (HF directory, or Meta's raw
``consolidated.00.pth`` + ``params.json``). Prints, per batch, the max / mean absolute logit
error and the loss difference, and at the end the mean of the per-batch maxima (the number the
reference's weight-conversion test gates on, tests/test_llama_weights.py:84-107).

The Meta baseline does not need the ``llama`` / ``fairscale`` packages: ``MetaLlama`` below is
a plain single-device PyTorch restatement of that architecture (interleaved-pair RoPE, RMSNorm,
SwiGLU, GQA) that consumes Meta's state-dict keys directly.
"""
import json
import math
import warnings
from pathlib import Path
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from epfl_megatron_b200 import get_args, update_num_microbatches
from epfl_megatron_b200.arguments import parse_args
from epfl_megatron_b200.initialize import initialize_megatron
from epfl_megatron_b200.models.enums import ModelType
from epfl_megatron_b200.training import build_train_valid_test_data_iterators, setup_model_and_optimizer
from finetune import extra_args, get_batch, loss_func, model_provider, train_valid_test_datasets_provider


class MetaLlama(nn.Module):
    """Reference-free restatement of Meta's Llama / Llama-2 decoder for parity checks."""

    def __init__(self, dim, n_layers, n_heads, n_kv_heads=None, vocab_size=32000, multiple_of=256,
                 ffn_dim_multiplier=None, norm_eps=1e-5, max_seq_len=4096, rope_theta=10000.0, **_):
        super().__init__()
        self.dim, self.n_layers, self.n_heads = dim, n_layers, n_heads
        self.n_kv_heads = n_kv_heads or n_heads
        self.eps, self.hd = norm_eps, dim // n_heads
        hidden = int(2 * (4 * dim) / 3)
        if ffn_dim_multiplier is not None:
            hidden = int(ffn_dim_multiplier * hidden)
        hidden = multiple_of * ((hidden + multiple_of - 1) // multiple_of)
        self.p = nn.ParameterDict()

        def add(name, *shape):
            self.p[name.replace(".", "/")] = nn.Parameter(torch.zeros(*shape), requires_grad=False)
        add("tok_embeddings.weight", vocab_size, dim)
        add("norm.weight", dim)
        add("output.weight", vocab_size, dim)
        for i in range(n_layers):
            add(f"layers.{i}.attention.wq.weight", n_heads * self.hd, dim)
            add(f"layers.{i}.attention.wk.weight", self.n_kv_heads * self.hd, dim)
            add(f"layers.{i}.attention.wv.weight", self.n_kv_heads * self.hd, dim)
            add(f"layers.{i}.attention.wo.weight", dim, n_heads * self.hd)
            add(f"layers.{i}.feed_forward.w1.weight", hidden, dim)
            add(f"layers.{i}.feed_forward.w2.weight", dim, hidden)
            add(f"layers.{i}.feed_forward.w3.weight", hidden, dim)
            add(f"layers.{i}.attention_norm.weight", dim)
            add(f"layers.{i}.ffn_norm.weight", dim)
        inv = 1.0 / (rope_theta ** (torch.arange(0, self.hd, 2).float() / self.hd))
        ang = torch.outer(torch.arange(max_seq_len).float(), inv)
        self.register_buffer("cos", ang.cos(), persistent=False)
        self.register_buffer("sin", ang.sin(), persistent=False)

    def load_meta_state_dict(self, sd):
        for k, v in sd.items():
            key = k.replace(".", "/")
            if key in self.p:
                self.p[key].data.copy_(v)

    def _w(self, name):
        return self.p[name.replace(".", "/")]

    def _norm(self, x, w):
        xf = x.float()
        return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.eps)).type_as(x) * w

    def _rope(self, x):
        b, s, h, d = x.shape
        xf = x.float().reshape(b, s, h, d // 2, 2)
        c, si = self.cos[:s, None, :], self.sin[:s, None, :]
        out = torch.stack([xf[..., 0] * c - xf[..., 1] * si, xf[..., 0] * si + xf[..., 1] * c], dim=-1)
        return out.reshape(b, s, h, d).type_as(x)

    def forward(self, input_ids, position_ids=None, attention_mask=None, labels=None):
        b, s = input_ids.shape
        h = F.embedding(input_ids, self._w("tok_embeddings.weight"))
        mask = torch.full((s, s), float("-inf"), device=h.device).triu(1)
        rep = self.n_heads // self.n_kv_heads
        for i in range(self.n_layers):
            pre = f"layers.{i}."
            x = self._norm(h, self._w(pre + "attention_norm.weight"))
            q = F.linear(x, self._w(pre + "attention.wq.weight")).view(b, s, self.n_heads, self.hd)
            k = F.linear(x, self._w(pre + "attention.wk.weight")).view(b, s, self.n_kv_heads, self.hd)
            v = F.linear(x, self._w(pre + "attention.wv.weight")).view(b, s, self.n_kv_heads, self.hd)
            q, k = self._rope(q), self._rope(k)
            k, v = k.repeat_interleave(rep, dim=2), v.repeat_interleave(rep, dim=2)
            sc = torch.einsum("bqhd,bkhd->bhqk", q, k) / math.sqrt(self.hd) + mask
            o = torch.einsum("bhqk,bkhd->bqhd", sc.float().softmax(-1).type_as(q), v).reshape(b, s, -1)
            h = h + F.linear(o, self._w(pre + "attention.wo.weight"))
            x = self._norm(h, self._w(pre + "ffn_norm.weight"))
            ff = F.silu(F.linear(x, self._w(pre + "feed_forward.w1.weight"))) * \
                F.linear(x, self._w(pre + "feed_forward.w3.weight"))
            h = h + F.linear(ff, self._w(pre + "feed_forward.w2.weight"))
        logits = F.linear(self._norm(h, self._w("norm.weight")), self._w("output.weight")).float()
        loss = torch.zeros((), device=logits.device)
        if labels is not None:
            warnings.warn("the Meta-format baseline does not compute a loss")
        return {"logits": logits, "loss": loss}


def is_meta_llama2_path(path: Optional[Path]) -> bool:
    return path is not None and len(list(Path(path).glob("*.pth"))) > 0


def meta_provider(cache_dir: Path):
    cache_dir = Path(cache_dir)
    ckpts = sorted(cache_dir.glob("*.pth"))
    assert len(ckpts) == 1, "only unsharded Meta checkpoints (one consolidated.00.pth) are supported"
    with open(cache_dir / "params.json") as f:
        params = json.load(f)
    if params.get("vocab_size", -1) <= 0:
        params["vocab_size"] = 32000
    model = MetaLlama(**params)
    model.load_meta_state_dict(torch.load(ckpts[0], map_location="cpu"))
    return model


def hf_provider(name: str, cache_dir: Optional[Path], device: str, size: int = 7):
    from transformers import AutoModelForCausalLM, LlamaForCausalLM
    print("Getting the baseline model...")
    if name == "falcon":
        src = cache_dir if cache_dir is not None and (Path(cache_dir) / "config.json").exists() \
            else f"tiiuae/falcon-{size}b"
        model = AutoModelForCausalLM.from_pretrained(src, cache_dir=None if src == cache_dir else cache_dir)
    elif name in ("llama", "llama2", "codellama") and is_meta_llama2_path(cache_dir):
        print(f"baseline path {cache_dir} holds raw Meta weights")
        model = meta_provider(cache_dir)
    elif name in ("llama", "codellama"):
        model = LlamaForCausalLM.from_pretrained(cache_dir)
    elif name == "llama2":
        src = cache_dir if cache_dir is not None and (Path(cache_dir) / "config.json").exists() \
            else f"meta-llama/Llama-2-{size}b-hf"
        model = LlamaForCausalLM.from_pretrained(src, cache_dir=None if src == cache_dir else cache_dir)
    else:
        raise KeyError(f"Model {name} not implemented")
    return model.eval().requires_grad_(False).to(device)


def hf_our_provider(name: str, data_dir: Path, device: str, size: int = 7):
    from transformers import AutoModelForCausalLM
    return AutoModelForCausalLM.from_pretrained(data_dir).eval().requires_grad_(False).to(device)


def hf_forward(model, batch):
    device = next(p.device for p in model.parameters())
    tokens, labels, loss_mask, attention_mask, position_ids = [
        t.to(device) if torch.is_tensor(t) else t for t in batch]
    out = model(input_ids=tokens, position_ids=position_ids, labels=tokens)
    return out["logits"], out["loss"]


def mega_provider(name: str):
    print("Getting the model of this framework...")
    model, _, _ = setup_model_and_optimizer(model_provider, ModelType.encoder_or_decoder)
    assert len(model) == 1, "correctness verification only supports unsharded models"
    return model[0].eval().requires_grad_(False)


def mega_forward(model, batch):
    tokens, labels, loss_mask, attention_mask, position_ids = batch
    assert torch.all(loss_mask.bool())
    logits = model(tokens, position_ids, attention_mask, labels=None)
    losses = model(tokens, position_ids, attention_mask, labels=labels)
    loss, _ = loss_func(loss_mask, losses)
    return logits, loss


def verify_step(our_forward, our_model, base_forward, base_model, batch):
    our_logits, our_loss = our_forward(our_model, batch)
    base_logits, base_loss = base_forward(base_model, batch)
    v = min(our_logits.size(-1), base_logits.size(-1))     # padded vocabulary rows are not compared
    our_logits, base_logits = our_logits[..., :v].float().cpu(), base_logits[..., :v].float().cpu()
    assert our_logits.size() == base_logits.size(), f"ours={our_logits.size()}, true={base_logits.size()}"
    err = (our_logits - base_logits).abs()
    print(f"Max absoulute error in the logits: max={err.max():.6f}, avg={err.mean():.6f}")
    our_loss, base_loss = our_loss.float().cpu(), base_loss.float().cpu()
    print(f"Abs loss error: {(our_loss - base_loss).abs():.6f} Our loss: {our_loss:.3f}, theirs: {base_loss:.3f}")
    return float(err.max()), float(err.mean())


def is_megatron_path(path):
    return path is not None and (Path(path) / "latest_checkpointed_iteration.txt").exists()


def main():
    print("Starting verification against the baseline")
    args = get_args()
    dev = "cuda:0" if torch.cuda.is_available() else "cpu"
    if is_megatron_path(args.load):
        our_model, our_forward = mega_provider(args.model_name), mega_forward
    else:
        print(f"NOTE: {args.load} does not look like a megatron checkpoint, assuming a huggingface one")
        our_model, our_forward = hf_our_provider(args.model_name, args.load, dev), hf_forward
        args.iteration = 0
    base_device = args.baseline_device if torch.cuda.is_available() else "cpu"
    base_model = hf_provider(args.model_name, args.cache_dir, base_device, size=args.model_size)
    data_iterator, _, _ = build_train_valid_test_data_iterators(train_valid_test_datasets_provider)
    maxima = []
    for iteration in range(args.verify_batches):
        print(f"Iteration {iteration}...")
        update_num_microbatches(args.consumed_train_samples)
        args.curr_iteration = iteration
        mx, _ = verify_step(our_forward, our_model, hf_forward, base_model, get_batch(data_iterator))
        maxima.append(mx)
    print(f"Mean of the per-batch max abs logit errors: {sum(maxima) / len(maxima):.6f}")


def extra_extra_args(parser):
    parser = extra_args(parser)
    group = parser.add_argument_group(title="huggingface")
    group.add_argument("--huggingface_cache", type=Path, default=None, dest="cache_dir",
                       help="baseline: HF checkpoint directory / cache, or Meta's raw weight directory")
    group.add_argument("--huggingface_device", default="cuda:1", dest="baseline_device")
    group.add_argument("--model_size", type=int, default=7)
    group.add_argument("--verify_batches", type=int, default=10)
    return parser


if __name__ == "__main__":
    defaults = {"micro_batch_size": 1, "use_checkpoint_args": True, "train_iters": 10, "lr": 1.0}
    if not is_megatron_path(parse_args(extra_extra_args).load):
        defaults.update({"encoder_num_layers": 1, "hidden_size": 1, "num_attention_heads": 1,
                         "seq_length": 2048, "max_position_embeddings": 2048})
    initialize_megatron(extra_extra_args, args_defaults=defaults)
    main()
