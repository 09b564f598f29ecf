"""Synthetic-weight round trip (stands in for the reference's tests/test_llama_weights.py,
which needs real Llama-2 weights): HF layout -> Megatron release checkpoint -> shard to
tp2/pp2 -> unshard -> HF layout must be bit-exact, and the Megatron model built from the
converted checkpoint must reproduce the HF model's logits (reference gate: <= 1e-3 in fp32)."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "weights2megatron"))


def _tiny_hf_llama(kv_heads=2):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=256, hidden_size=64, intermediate_size=176, num_hidden_layers=4,
                      num_attention_heads=4, num_key_value_heads=kv_heads, max_position_embeddings=64,
                      rms_norm_eps=1e-5, tie_word_embeddings=False)
    return LlamaForCausalLM(cfg).float().eval(), cfg


def _run(cmd):
    # offline box: make hub look-ups fail at once instead of after connection retries
    env = dict(os.environ, HF_HUB_OFFLINE="1", TRANSFORMERS_OFFLINE="1")
    r = subprocess.run([sys.executable] + cmd, capture_output=True, text=True, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r


@pytest.mark.parametrize("kv_heads", [2, 4])
def test_llama_roundtrip_bit_exact(tmp_path, kv_heads):
    import weights2megatron as w2m
    import megatron2hf as m2h
    hf, cfg = _tiny_hf_llama(kv_heads)
    mega = tmp_path / "mega"
    w2m.convert("llama2", 7, mega, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    assert (mega / "latest_checkpointed_iteration.txt").read_text() == "release"
    assert (mega / "release" / "mp_rank_00" / "model_optim_rng.pt").exists()
    sharded, back = tmp_path / "tp2pp2", tmp_path / "tp1pp1"
    _run(["tools/checkpoint_util.py", "--model_type", "llama2", "--load_dir", str(mega), "--save_dir", str(sharded),
          "--target_tensor_parallel_size", "2", "--target_pipeline_parallel_size", "2"])
    for t in range(2):
        for p in range(2):
            assert (sharded / "release" / f"mp_rank_{t:02d}_{p:03d}" / "model_optim_rng.pt").exists()
    sd = torch.load(sharded / "release" / "mp_rank_01_001" / "model_optim_rng.pt", weights_only=False)
    enc = sd["model"]["language_model"]["encoder"]
    assert enc["layers.0.self_attention.query_key_value.weight"].shape == (
        (cfg.num_attention_heads + 2 * kv_heads) * 16 // 2, 64)
    assert enc["layers.1.mlp.dense_h_to_4h.weight"].shape == (176, 64)      # [up_shard; gate_shard]
    assert "lm_head" in sd["model"]["language_model"] and "embedding" not in sd["model"]["language_model"]
    _run(["tools/checkpoint_util.py", "--model_type", "llama2", "--load_dir", str(sharded), "--save_dir", str(back),
          "--target_tensor_parallel_size", "1", "--target_pipeline_parallel_size", "1"])
    lm, args = m2h.load_megatron(back)
    out = m2h.llama_state_dict(lm, args)
    ref = hf.state_dict()
    for k, v in out.items():
        assert torch.equal(v, ref[k]), k
    # and the HF writer produces a loadable model with identical logits
    m2h.write_llama_model(tmp_path / "hf_out", back, dtype=torch.float32, safe=False)
    from transformers import LlamaForCausalLM
    again = LlamaForCausalLM.from_pretrained(tmp_path / "hf_out", torch_dtype=torch.float32).eval()
    ids = torch.randint(0, 256, (2, 32))
    with torch.no_grad():
        assert torch.allclose(again(ids).logits, hf(ids).logits, atol=1e-6)


def _megatron_logits(ckpt_dir, ids, tmp_path):
    """Build our LlamaModel from a converted checkpoint in a subprocess and return its logits."""
    out_file = tmp_path / "logits.pt"
    torch.save(ids, tmp_path / "ids.pt")
    code = f"""
import sys, os, torch
sys.path.insert(0, {ROOT!r})
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="{29000 + os.getpid() % 2000}", RANK="0", WORLD_SIZE="1")
import finetune
from epfl_megatron_b200 import get_args
from epfl_megatron_b200.initialize import initialize_megatron
from epfl_megatron_b200.training import get_model
from epfl_megatron_b200.checkpointing import load_checkpoint
from epfl_megatron_b200.models.enums import ModelType
argv = ["--model_name", "llama2", "--load", {str(ckpt_dir)!r}, "--use_checkpoint_args", "--micro_batch_size", "2",
        "--distributed_backend", "gloo", "--no_load_optim", "--no_load_rng", "--finetune", "--use_flash_attn",
        "--hidden_dropout", "0", "--attention_dropout", "0", "--no_bias_gelu_fusion", "--tokenizer_type", "NullTokenizer",
        "--train_iters", "1", "--lr", "1e-4"]
initialize_megatron(finetune.extra_args, {{}}, args_list=argv)
args = get_args()
model = get_model(finetune.model_provider, ModelType.encoder_or_decoder, wrap_with_ddp=False)
load_checkpoint(model, None, None)
model[0].eval()
ids = torch.load({str(tmp_path / 'ids.pt')!r})
pos = torch.arange(ids.shape[1]).unsqueeze(0).expand_as(ids)
with torch.no_grad():
    logits = model[0](ids, pos, None)
torch.save(logits.float(), {str(out_file)!r})
"""
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    return torch.load(out_file)


def test_converted_llama_reproduces_hf_logits(tmp_path):
    import weights2megatron as w2m
    hf, cfg = _tiny_hf_llama(2)
    mega = tmp_path / "mega"
    w2m.convert("llama2", 7, mega, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    ids = torch.randint(0, 256, (2, 48))
    ours = _megatron_logits(mega, ids, tmp_path)
    with torch.no_grad():
        ref = hf(ids).logits
    err = (ours - ref).abs().max().item()
    assert err <= 1e-3, f"max abs logit error {err}"


def test_falcon_roundtrip_bit_exact(tmp_path):
    import weights2megatron as w2m
    import megatron2hf as m2h
    from transformers import FalconConfig, FalconForCausalLM
    torch.manual_seed(1)
    cfg = FalconConfig(vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, num_kv_heads=2,
                       new_decoder_architecture=True, parallel_attn=True, bias=False, alibi=False)
    hf = FalconForCausalLM(cfg).float().eval()
    mega = tmp_path / "mega"
    w2m.convert("falcon", 40, mega, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    lm, args = m2h.load_megatron(mega)
    assert args.parallel_attn and args.parallel_layernorm and args.num_attention_heads_kv == 2
    out = m2h.falcon_state_dict(lm, args)
    ref = hf.state_dict()
    for k, v in out.items():
        assert torch.equal(v, ref[k]), k


def test_permute_qkv_is_an_involution_pair():
    from epfl_megatron_b200.utils.ckpt_convert import permute_qkv
    w = torch.randn((8 + 2 * 2) * 16, 128)
    there = permute_qkv(w, 128, 8, 2)
    assert not torch.equal(there, w)
    assert torch.equal(permute_qkv(there, 128, 8, 2, revert=True), w)
    hn, qpg = 16, 4
    g = w.view(2, qpg + 2, hn, 128)
    assert torch.equal(there.view(2, qpg + 2, hn, 128)[:, qpg + 1], g[:, qpg + 1])   # v untouched


def test_loader_saver_plugins_match_direct_reshard(tmp_path):
    """The queue protocol (loader plug-in -> saver plug-in) writes the same shards as the
    direct file-level path of tools/checkpoint_util.py."""
    import queue as pyqueue
    import types
    import weights2megatron as w2m
    sys.path.insert(0, ROOT)
    from tools import checkpoint_loader_megatron as loader, checkpoint_saver_megatron as saver
    hf, cfg = _tiny_hf_llama(2)
    mega = tmp_path / "mega"
    w2m.convert("llama2", 7, mega, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    direct, via_queue = tmp_path / "direct", tmp_path / "queue"
    _run(["tools/checkpoint_util.py", "--model_type", "llama2", "--load_dir", str(mega), "--save_dir", str(direct),
          "--target_tensor_parallel_size", "2", "--target_pipeline_parallel_size", "2"])
    q = pyqueue.Queue()
    a = types.SimpleNamespace(model_type="llama2", load_dir=str(mega), save_dir=str(via_queue), checking=True,
                              true_vocab_size=None, target_tensor_parallel_size=2,
                              target_pipeline_parallel_size=2)
    loader.load_checkpoint(q, a)
    saver.save_checkpoint(q, a)
    for t in range(2):
        for p in range(2):
            d = torch.load(direct / "release" / f"mp_rank_{t:02d}_{p:03d}" / "model_optim_rng.pt", weights_only=False)
            v = torch.load(via_queue / "release" / f"mp_rank_{t:02d}_{p:03d}" / "model_optim_rng.pt", weights_only=False)

            def flat(x, pre=""):
                out = {}
                for k, val in x.items():
                    if isinstance(val, dict):
                        out.update(flat(val, pre + k + "/"))
                    else:
                        out[pre + k] = val
                return out
            fd, fv = flat(d["model"]), flat(v["model"])
            assert fd.keys() == fv.keys()
            for k in fd:
                assert torch.equal(fd[k], fv[k]), k


def test_meta_shards_merge_and_convert_like_the_hf_route(tmp_path):
    """Meta's model-parallel ``consolidated.NN.pth`` shards are merged along Meta's per-key dimensions,
    and the Meta route (weights already in interleaved-pair RoPE order) produces the same Megatron
    tensors as the HF route (half-rotation order, permuted on the way in)."""
    import sys as _sys
    _sys.path.insert(0, os.path.join(ROOT, "weights2megatron"))
    try:
        import merge_llama as ml
        import weights2megatron as w2m
    finally:
        _sys.path.pop(0)
    from epfl_megatron_b200.utils.ckpt_convert import rope_hf_to_interleaved
    torch.manual_seed(3)
    n_layer, hidden, n_heads, n_kv, ffn, vocab = 2, 64, 4, 2, 96, 128
    hn = hidden // n_heads
    hf = {"model.embed_tokens.weight": torch.randn(vocab, hidden), "model.norm.weight": torch.randn(hidden),
          "lm_head.weight": torch.randn(vocab, hidden)}
    for i in range(n_layer):
        p = f"model.layers.{i}"
        hf.update({f"{p}.self_attn.q_proj.weight": torch.randn(n_heads * hn, hidden),
                   f"{p}.self_attn.k_proj.weight": torch.randn(n_kv * hn, hidden),
                   f"{p}.self_attn.v_proj.weight": torch.randn(n_kv * hn, hidden),
                   f"{p}.self_attn.o_proj.weight": torch.randn(hidden, hidden),
                   f"{p}.mlp.gate_proj.weight": torch.randn(ffn, hidden), f"{p}.mlp.up_proj.weight": torch.randn(ffn, hidden),
                   f"{p}.mlp.down_proj.weight": torch.randn(hidden, ffn),
                   f"{p}.input_layernorm.weight": torch.randn(hidden),
                   f"{p}.post_attention_layernorm.weight": torch.randn(hidden),
                   f"{p}.self_attn.rotary_emb.inv_freq": torch.randn(hn // 2)})
    named = ml.hf_state_to_meta(hf)
    assert "layers.1.feed_forward.w3.weight" in named and not any("rotary" in k for k in named)
    # the same model as Meta would publish it: q/k rows in interleaved-pair order
    meta = dict(named)
    for i in range(n_layer):
        meta[f"layers.{i}.attention.wq.weight"] = rope_hf_to_interleaved(named[f"layers.{i}.attention.wq.weight"], n_heads, hn)
        meta[f"layers.{i}.attention.wk.weight"] = rope_hf_to_interleaved(named[f"layers.{i}.attention.wk.weight"], n_kv, hn)
    meta["rope.freqs"] = torch.randn(hn // 2)
    # cut into two model-parallel shards the way Meta does
    col = ("wq", "wk", "wv", "w1", "w3", "output")              # split along dim 0
    row = ("wo", "w2", "tok_embeddings")                       # split along the last dim
    shards = [{}, {}]
    for k, v in meta.items():
        short = k.split(".")[-2] if "." in k else k
        for r in range(2):
            if short in col:
                shards[r][k] = v.chunk(2, dim=0)[r].clone()
            elif short in row:
                shards[r][k] = v.chunk(2, dim=-1)[r].clone()
            else:
                shards[r][k] = v.clone()
    for r in range(2):
        torch.save(shards[r], tmp_path / f"consolidated.{r:02d}.pth")
    merged, source = ml.merge_llama(7, 2, root_dir=str(tmp_path))
    assert source == "meta" and set(merged) == set(meta)
    for k in meta:
        assert torch.equal(merged[k], meta[k]), k
    via_meta = w2m.llama_to_megatron(merged, n_layer, hidden, n_heads, n_kv, source="meta")
    via_hf = w2m.llama_to_megatron(named, n_layer, hidden, n_heads, n_kv, source="hf")
    for k in via_hf["encoder"]:
        assert torch.equal(via_meta["encoder"][k], via_hf["encoder"][k]), k
    assert torch.equal(via_meta["lm_head"], hf["lm_head.weight"])
    glu = via_hf["encoder"]["layers.0.mlp.dense_h_to_4h.weight"]
    assert torch.equal(glu[:ffn], hf["model.layers.0.mlp.up_proj.weight"])      # [up; gate]


def test_meta_llama_to_hf_cli(tmp_path):
    """weights2megatron/convert_llama2hf.py: two Meta shards + params.json -> a HF directory whose
    weights equal the HF model the shards were derived from."""
    transformers = pytest.importorskip("transformers")
    from epfl_megatron_b200.utils.ckpt_convert import rope_hf_to_interleaved
    torch.manual_seed(5)
    cfg = transformers.LlamaConfig(hidden_size=64, intermediate_size=96, num_attention_heads=4, num_key_value_heads=2,
                                   num_hidden_layers=2, vocab_size=128, rms_norm_eps=1e-5, tie_word_embeddings=False)
    hf = transformers.LlamaForCausalLM(cfg).float()
    sd = {k: v.detach().clone() for k, v in hf.state_dict().items() if "rotary" not in k}
    hn = 16
    meta = {"tok_embeddings.weight": sd["model.embed_tokens.weight"], "norm.weight": sd["model.norm.weight"],
            "output.weight": sd["lm_head.weight"], "rope.freqs": torch.zeros(hn // 2)}
    names = {"self_attn.q_proj": "attention.wq", "self_attn.k_proj": "attention.wk", "self_attn.v_proj": "attention.wv",
             "self_attn.o_proj": "attention.wo", "mlp.gate_proj": "feed_forward.w1", "mlp.down_proj": "feed_forward.w2",
             "mlp.up_proj": "feed_forward.w3", "input_layernorm": "attention_norm",
             "post_attention_layernorm": "ffn_norm"}
    for i in range(2):
        for h, m in names.items():
            w = sd[f"model.layers.{i}.{h}.weight"]
            if m == "attention.wq":
                w = rope_hf_to_interleaved(w, 4, hn)
            elif m == "attention.wk":
                w = rope_hf_to_interleaved(w, 2, hn)
            meta[f"layers.{i}.{m}.weight"] = w
    src = tmp_path / "meta"
    src.mkdir()
    col, row = ("wq", "wk", "wv", "w1", "w3", "output"), ("wo", "w2", "tok_embeddings")
    for r in range(2):
        shard = {}
        for k, v in meta.items():
            short = k.split(".")[-2] if "." in k else k
            shard[k] = (v.chunk(2, dim=0)[r] if short in col else v.chunk(2, dim=-1)[r] if short in row else v).clone()
        torch.save(shard, src / f"consolidated.{r:02d}.pth")
    (src / "params.json").write_text(json.dumps({"dim": 64, "n_layers": 2, "n_heads": 4, "n_kv_heads": 2,
                                                 "norm_eps": 1e-5, "vocab_size": -1}))
    spm = pytest.importorskip("sentencepiece")
    (tmp_path / "c.txt").write_text("\n".join(["the quick brown fox jumps over the lazy dog"] * 60))
    spm.SentencePieceTrainer.train(input=str(tmp_path / "c.txt"), model_prefix=str(src / "tokenizer"), vocab_size=40,
                                   model_type="bpe", bos_id=1, eos_id=2, unk_id=0, pad_id=-1)
    out = tmp_path / "hf_out"
    _run(["weights2megatron/convert_llama2hf.py", "--input_dir", str(src), "--model_size", "13",
          "--output_dir", str(out)])
    tok = transformers.AutoTokenizer.from_pretrained(str(out))
    assert len(tok) == 40 and len(tok("the quick fox")["input_ids"]) >= 3
    back = transformers.LlamaForCausalLM.from_pretrained(str(out), torch_dtype=torch.float32)
    got = back.state_dict()
    for k, v in sd.items():
        assert torch.equal(got[k], v), k
    assert back.config.num_key_value_heads == 2 and back.config.intermediate_size == 96


def test_verify_correctness_cli(tmp_path):
    """verify_correctness.py as a user runs it: a converted Megatron checkpoint against the HF model it
    came from, on batches drawn from a tokenised corpus; reports max / mean absolute logit error per batch
    and the loss difference (reference: verify_correctness.py:107-166, gate 1e-3 in tests/test_llama_weights.py)."""
    import re
    import numpy as np
    import weights2megatron as w2m
    from epfl_megatron_b200.data import indexed_dataset as ids
    from tests.helpers.launch import free_port
    hf, cfg = _tiny_hf_llama(2)
    hf.save_pretrained(tmp_path / "hf")
    w2m.convert("llama2", 7, tmp_path / "mega", dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    rng = np.random.default_rng(0)
    b = ids.MMapIndexedDatasetBuilder(ids.data_file_path(str(tmp_path / "corpus")), dtype=np.uint16)
    for _ in range(100):
        b.add_item(torch.tensor(rng.integers(1, 250, size=int(rng.integers(30, 120)))))
        b.end_document()
    b.finalize(ids.index_file_path(str(tmp_path / "corpus")))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    cmd = [sys.executable, os.path.join(ROOT, "verify_correctness.py"), "--model_name", "llama2",
           "--load", str(tmp_path / "mega"), "--data_path", str(tmp_path / "corpus"), "--split", "100,0,0",
           "--tokenizer_type", "NullTokenizer", "--vocab_size", "256", "--make_vocab_size_divisible_by", "1",
           "--huggingface_cache", str(tmp_path / "hf"), "--verify_batches", "3", "--distributed_backend", "gloo",
           "--use_cpu_initialization", "--no_load_optim", "--no_load_rng", "--num_workers", "0", "--eval_iters", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-2500:] + r.stderr[-2500:]
    maxima = [float(x) for x in re.findall(r"Max absoulute error in the logits: max=([0-9.]+)", r.stdout)]
    assert len(maxima) == 3 and max(maxima) <= 1e-3, r.stdout[-1500:]
    mean = float(re.search(r"Mean of the per-batch max abs logit errors: ([0-9.]+)", r.stdout).group(1))
    assert mean <= 1e-3
    diffs = [float(x) for x in re.findall(r"Abs loss error: ([0-9.]+)", r.stdout)]
    # the HF side scores positions 0..s-2 of the same tokens (labels shifted inside the model), this
    # side all s positions against the data's labels: equal logits, slightly different mean loss
    assert len(diffs) == 3 and max(diffs) <= 2e-2


def test_megatron2hf_cli_with_tokenizer(tmp_path):
    """weights2megatron/megatron2hf.py from the command line: the exported HF directory loads with
    transformers, reproduces the source weights, and carries a tokenizer whose added special tokens sit
    at the ids the Megatron SentencePiece tokenizer gave them."""
    spm = pytest.importorskip("sentencepiece")
    transformers = pytest.importorskip("transformers")
    import weights2megatron as w2m
    hf, cfg = _tiny_hf_llama(2)
    mega = tmp_path / "mega"
    w2m.convert("llama2", 7, mega, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    corpus = tmp_path / "c.txt"
    corpus.write_text("\n".join(["the quick brown fox jumps over the lazy dog", "pack my box with five dozen jugs"] * 40))
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=str(tmp_path / "tok"), vocab_size=60,
                                   model_type="bpe", bos_id=1, eos_id=2, unk_id=0, pad_id=-1)
    # the release checkpoint's args name the tokenizer that was used for training
    ck = mega / "release" / "mp_rank_00" / "model_optim_rng.pt"
    sd = torch.load(ck, weights_only=False)
    sd["args"].tokenizer_type, sd["args"].vocab_extra_ids, sd["args"].vocab_extra_ids_list = "SentencePieceTokenizer", 0, None
    sd["args"].new_tokens = True
    torch.save(sd, ck)
    out = tmp_path / "hf_out"
    _run(["weights2megatron/megatron2hf.py", "--model", "llama2", "--input_dir", str(mega), "--output_dir", str(out),
          "--vocab_file", str(tmp_path / "tok.model"), "--dtype", "fp32"])
    back = transformers.LlamaForCausalLM.from_pretrained(str(out), torch_dtype=torch.float32)
    for k, v in hf.state_dict().items():
        if "rotary" not in k:
            assert torch.equal(back.state_dict()[k], v), k
    tok = transformers.AutoTokenizer.from_pretrained(str(out))
    from epfl_megatron_b200.tokenizer.tokenizer import _SentencePieceTokenizer
    mt = _SentencePieceTokenizer(str(tmp_path / "tok.model"), new_tokens=True)
    for name in ("<CLS>", "<SEP>", "<EOD>", "<MASK>"):
        assert tok.convert_tokens_to_ids(name) == mt.vocab[name], name
    assert len(tok) == mt.vocab_size


def test_weights2megatron_cli_reads_dimensions_from_the_source(tmp_path):
    """weights2megatron.py from the command line on a HF directory of a non-standard size: the model
    dimensions come from its config.json (the built-in size table only knows the published checkpoints),
    and the release checkpoint equals the in-memory conversion."""
    import weights2megatron as w2m
    hf, cfg = _tiny_hf_llama(2)
    hf.save_pretrained(tmp_path / "hf")
    _run(["weights2megatron/weights2megatron.py", "llama2", "--size", "7", "--out", str(tmp_path / "cli"),
          "--model_path", str(tmp_path / "hf"), "--dtype", "fp32"])
    w2m.convert("llama2", 7, tmp_path / "api", dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    a = torch.load(tmp_path / "cli" / "release" / "mp_rank_00" / "model_optim_rng.pt", weights_only=False)
    b = torch.load(tmp_path / "api" / "release" / "mp_rank_00" / "model_optim_rng.pt", weights_only=False)
    assert (tmp_path / "cli" / "latest_checkpointed_iteration.txt").read_text() == "release"
    for k in ("num_layers", "hidden_size", "num_attention_heads", "num_attention_heads_kv", "ffn_hidden_size",
              "max_position_embeddings", "padded_vocab_size"):
        assert getattr(a["args"], k) == getattr(b["args"], k), k
    assert a["args"].num_layers == 4 and a["args"].ffn_hidden_size == 176

    def flat(d, prefix=""):
        for k, v in d.items():
            if isinstance(v, dict):
                yield from flat(v, prefix + k + ".")
            else:
                yield prefix + k, v
    fa, fb = dict(flat(a["model"]["language_model"])), dict(flat(b["model"]["language_model"]))
    assert fa.keys() == fb.keys()
    for k in fa:
        assert torch.equal(fa[k], fb[k]), k


def test_permute_qkv_migration_cli(tmp_path):
    """weights2megatron/permute_qkv.py: a checkpoint whose fused QKV rows were left in HF's half-rotation
    order (converted without the permutation) becomes identical to a correctly converted one."""
    import weights2megatron as w2m
    hf, cfg = _tiny_hf_llama(2)
    good = tmp_path / "good"
    w2m.convert("llama2", 7, good, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    # the faulty conversion: same pipeline, but the q/k rows are taken as if they were already interleaved
    import merge_llama as ml
    lm = w2m.llama_to_megatron(ml.hf_state_to_meta(hf.state_dict()), 4, 64, 4, 2, source="meta")
    ck = torch.load(good / "release" / "mp_rank_00" / "model_optim_rng.pt", weights_only=False)
    bad = tmp_path / "bad"
    w2m.save_release(lm, ck["args"], bad, torch.float32)
    fixed = tmp_path / "fixed"
    _run(["weights2megatron/permute_qkv.py", "--input_dir", str(bad), "--output_dir", str(fixed)])
    a = torch.load(fixed / "release" / "mp_rank_00" / "model_optim_rng.pt", weights_only=False)
    ea, eg = a["model"]["language_model"]["encoder"], ck["model"]["language_model"]["encoder"]
    bad_enc = torch.load(bad / "release" / "mp_rank_00" / "model_optim_rng.pt",
                         weights_only=False)["model"]["language_model"]["encoder"]
    key = "layers.0.self_attention.query_key_value.weight"
    assert not torch.equal(bad_enc[key], eg[key])
    for k in eg:
        assert torch.equal(ea[k], eg[k]), k
    assert (fixed / "latest_checkpointed_iteration.txt").read_text() == "release"


def test_falcon_cli_round_trip(tmp_path):
    """Falcon through the command lines in both directions: HF directory -> weights2megatron.py ->
    megatron2hf.py -> HF directory with identical weights and logits (40B-style block: grouped KV heads,
    two LayerNorms)."""
    from transformers import FalconConfig, FalconForCausalLM
    torch.manual_seed(2)
    cfg = FalconConfig(vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, num_kv_heads=2,
                       new_decoder_architecture=True, parallel_attn=True, bias=False, alibi=False)
    hf = FalconForCausalLM(cfg).float().eval()
    hf.save_pretrained(tmp_path / "hf")
    _run(["weights2megatron/weights2megatron.py", "falcon", "--size", "40", "--out", str(tmp_path / "mega"),
          "--model_path", str(tmp_path / "hf"), "--dtype", "fp32"])
    ck = torch.load(tmp_path / "mega" / "release" / "mp_rank_00" / "model_optim_rng.pt", weights_only=False)
    assert ck["args"].parallel_layernorm and ck["args"].num_attention_heads_kv == 2
    _run(["weights2megatron/megatron2hf.py", "--model", "falcon", "--input_dir", str(tmp_path / "mega"),
          "--output_dir", str(tmp_path / "back"), "--dtype", "fp32"])
    back = FalconForCausalLM.from_pretrained(str(tmp_path / "back"), torch_dtype=torch.float32).eval()
    ref = hf.state_dict()
    for k, v in back.state_dict().items():
        assert torch.equal(v, ref[k]), k
    ids = torch.randint(0, 128, (2, 24))
    with torch.no_grad():
        assert torch.equal(back(ids).logits, hf(ids).logits)


def test_falcon_multi_query_block_round_trip(tmp_path):
    """The 7B-style Falcon block (one KV head, one LayerNorm) through the conversion functions."""
    import weights2megatron as w2m
    import megatron2hf as m2h
    from transformers import FalconConfig, FalconForCausalLM
    torch.manual_seed(4)
    cfg = FalconConfig(vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, multi_query=True,
                       new_decoder_architecture=False, parallel_attn=True, bias=False, alibi=False)
    hf = FalconForCausalLM(cfg).float().eval()
    w2m.convert("falcon", 7, tmp_path / "mega", dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    lm, args = m2h.load_megatron(tmp_path / "mega")
    assert args.parallel_attn and not args.parallel_layernorm and args.num_attention_heads_kv == 1
    ref = hf.state_dict()
    for k, v in m2h.falcon_state_dict(lm, args).items():
        assert torch.equal(v, ref[k]), k


def test_reshard_repads_the_vocabulary(tmp_path):
    """checkpoint_util --true_vocab_size: rows beyond the real vocabulary are dropped and the table is
    padded again (with copies of the last real row) for the target tensor-parallel size (multiple of make_vocab_size_divisible_by x tp), in
    the embedding and the untied LM head alike (tools/checkpoint_saver_megatron.py:190-226)."""
    import weights2megatron as w2m
    hf, cfg = _tiny_hf_llama(4)                       # vocab 256, 4 KV heads (shardable 4 ways)
    mega = tmp_path / "mega"
    w2m.convert("llama2", 7, mega, dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    ck = mega / "release" / "mp_rank_00" / "model_optim_rng.pt"
    sd = torch.load(ck, weights_only=False)
    sd["args"].make_vocab_size_divisible_by = 64
    torch.save(sd, ck)
    out = tmp_path / "tp2"
    _run(["tools/checkpoint_util.py", "--model_type", "llama2", "--load_dir", str(mega), "--save_dir", str(out),
          "--target_tensor_parallel_size", "2", "--target_pipeline_parallel_size", "1", "--true_vocab_size", "200"])
    shards = [torch.load(out / "release" / f"mp_rank_0{t}" / "model_optim_rng.pt", weights_only=False) for t in range(2)]
    assert shards[0]["args"].padded_vocab_size == 256 and shards[0]["args"].tensor_model_parallel_size == 2
    emb = torch.cat([s["model"]["language_model"]["embedding"]["word_embeddings"]["weight"] for s in shards])
    head = torch.cat([s["model"]["language_model"]["lm_head"] for s in shards])
    ref_e, ref_h = hf.state_dict()["model.embed_tokens.weight"], hf.state_dict()["lm_head.weight"]
    assert emb.shape == (256, 64) and torch.equal(emb[:200], ref_e[:200]) and torch.equal(head[:200], ref_h[:200])
    # the padding rows replicate the last real row, as the reference's saver does
    assert torch.equal(emb[200:], ref_e[199:200].expand(56, -1)) and torch.equal(head[200:], ref_h[199:200].expand(56, -1))
    # a target that needs a larger table: 200 real rows -> multiple of 64 x 4
    out4 = tmp_path / "tp4"
    _run(["tools/checkpoint_util.py", "--model_type", "llama2", "--load_dir", str(mega), "--save_dir", str(out4),
          "--target_tensor_parallel_size", "4", "--target_pipeline_parallel_size", "1", "--true_vocab_size", "200"])
    s0 = torch.load(out4 / "release" / "mp_rank_00" / "model_optim_rng.pt", weights_only=False)
    assert s0["args"].padded_vocab_size == 256
    assert s0["model"]["language_model"]["embedding"]["word_embeddings"]["weight"].shape == (64, 64)


@pytest.mark.parametrize("new_arch,kv_heads", [(True, 2), (False, 1)])
def test_verify_correctness_cli_falcon(tmp_path, new_arch, kv_heads):
    """verify_correctness.py --model_name falcon for both Falcon layouts (40B-style grouped KV with two
    layer norms, 7B-style multi-query with one): converted weights reproduce the HF logits."""
    import re
    import numpy as np
    import weights2megatron as w2m
    from transformers import FalconConfig, FalconForCausalLM
    from epfl_megatron_b200.data import indexed_dataset as ids
    from tests.helpers.launch import free_port
    torch.manual_seed(0)
    cfg = FalconConfig(vocab_size=256, hidden_size=64, num_hidden_layers=2, num_attention_heads=4,
                       num_kv_heads=kv_heads, new_decoder_architecture=new_arch, multi_query=not new_arch,
                       parallel_attn=True, bias=False, alibi=False)
    hf = FalconForCausalLM(cfg).float().eval()
    hf.save_pretrained(tmp_path / "hf")
    w2m.convert("falcon", 7, tmp_path / "mega", dtype=torch.float32, state_dict=hf.state_dict(), config=cfg)
    rng = np.random.default_rng(0)
    b = ids.MMapIndexedDatasetBuilder(ids.data_file_path(str(tmp_path / "corpus")), dtype=np.uint16)
    for _ in range(100):
        b.add_item(torch.tensor(rng.integers(1, 250, size=int(rng.integers(30, 120)))))
        b.end_document()
    b.finalize(ids.index_file_path(str(tmp_path / "corpus")))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    cmd = [sys.executable, os.path.join(ROOT, "verify_correctness.py"), "--model_name", "falcon",
           "--load", str(tmp_path / "mega"), "--data_path", str(tmp_path / "corpus"), "--split", "100,0,0",
           "--tokenizer_type", "NullTokenizer", "--vocab_size", "256", "--make_vocab_size_divisible_by", "1",
           "--huggingface_cache", str(tmp_path / "hf"), "--verify_batches", "2", "--distributed_backend", "gloo",
           "--use_cpu_initialization", "--no_load_optim", "--no_load_rng", "--num_workers", "0", "--eval_iters", "0",
           "--seq_length", "64", "--max_position_embeddings", "64"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-2500:] + r.stderr[-2500:]
    maxima = [float(x) for x in re.findall(r"Max absoulute error in the logits: max=([0-9.]+)", r.stdout)]
    assert len(maxima) == 2 and max(maxima) <= 1e-3, r.stdout[-1500:]
