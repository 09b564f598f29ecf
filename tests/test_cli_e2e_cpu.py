"""The user-facing path end to end on CPU / Gloo: a tokenised corpus on disk, ``finetune.py`` from
the command line (pretrain(): data iterators, train loop, periodic evaluation, logging, checkpoint),
then a second invocation that resumes from the checkpoint and finishes the run.
Reference: finetune.py:141-150 -> megatron/training.py:54-167, :639-807, :855-939."""
import json
import math
import os
import re
import subprocess
import sys

import numpy as np
import torch

from epfl_megatron_b200.data import indexed_dataset as ids

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _corpus(prefix):
    rng = np.random.default_rng(0)
    b = ids.MMapIndexedDatasetBuilder(ids.data_file_path(prefix), dtype=np.uint16)
    for _ in range(300):
        b.add_item(torch.tensor(rng.integers(1, 1000, size=int(rng.integers(20, 200)))))
        b.end_document()
    b.finalize(ids.index_file_path(prefix))


def _finetune(tmp_path, port, extra):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(port), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    cmd = [sys.executable, os.path.join(ROOT, "finetune.py"), "--model_name", "llama2",
           "--num_layers", "2", "--hidden_size", "64", "--num_attention_heads", "4", "--num_attention_heads_kv", "2",
           "--ffn_hidden_size", "128", "--seq_length", "32", "--max_position_embeddings", "32",
           "--micro_batch_size", "2", "--global_batch_size", "4", "--lr", "1e-3", "--min_lr", "1e-4",
           "--lr_decay_style", "cosine", "--lr_warmup_iters", "2", "--lr_decay_iters", "9", "--use_rms_norm", "--glu_activation", "swiglu",
           "--no_tie_embed_logits", "--position_embedding_type", "rotary", "--no_bias_gelu_fusion",
           "--hidden_dropout", "0", "--attention_dropout", "0", "--tokenizer_type", "NullTokenizer",
           "--vocab_size", "1024", "--data_path", str(tmp_path / "corpus"), "--split", "90,5,5",
           "--distributed_backend", "gloo", "--log_interval", "2", "--eval_interval", "3", "--eval_iters", "2",
           "--save", str(tmp_path / "ckpt"), "--save_interval", "3", "--num_workers", "0",
           "--use_cpu_initialization", "--log_throughput", "--metrics_jsonl", str(tmp_path / "m.jsonl")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_finetune_cli_train_eval_save_resume(tmp_path):
    from tests.helpers.launch import free_port
    _corpus(str(tmp_path / "corpus"))
    out = _finetune(tmp_path, free_port(), ["--train_iters", "6"])
    its = [int(m) for m in re.findall(r"iteration\s+(\d+)/\s*6", out)]
    assert its == [2, 4, 6], out[-2000:]
    assert "lm loss" in out and "tokens/s" in out and "validation loss at iteration 3" in out
    assert open(tmp_path / "ckpt" / "latest_checkpointed_iteration.txt").read().strip() == "6"
    assert os.path.isfile(tmp_path / "ckpt" / "iter_0000003" / "mp_rank_00" / "model_optim_rng.pt")
    rows = [json.loads(l) for l in open(tmp_path / "m.jsonl")]
    assert any(r.get("name") == "lm loss" for r in rows) or any("tokens_per_s" in r for r in rows), rows[:3]
    # the sample-index caches were written next to the corpus
    assert any(f.endswith(".npy") for f in os.listdir(tmp_path))
    # second invocation: loads iteration 6 and runs to 9. A longer run changes the schedule lengths
    # derived from --train_iters; like the reference this is refused unless the checkpoint's schedule
    # is adopted (--use_checkpoint_opt_param_scheduler) or overridden.
    out2 = _finetune(tmp_path, free_port(), ["--train_iters", "9", "--load", str(tmp_path / "ckpt"),
                                             "--use_checkpoint_opt_param_scheduler"])
    assert "successfully loaded checkpoint" in out2 and "at iteration 6" in out2, out2[-2000:]
    its2 = [int(m) for m in re.findall(r"iteration\s+(\d+)/\s*9", out2)]
    assert its2 and its2[0] > 6 and its2[-1] <= 9
    assert open(tmp_path / "ckpt" / "latest_checkpointed_iteration.txt").read().strip() == "9"


def test_sigterm_checkpoints_and_exits(tmp_path):
    """--exit_signal_handler: a SIGTERM (here injected into rank 0 at iteration 3) makes the run save a
    checkpoint at that iteration and stop; --exit_interval stops a run at a multiple of the interval
    (megatron/training.py:712-748, dist_signal_handler.py)."""
    from tests.helpers.launch import free_port
    _corpus(str(tmp_path / "corpus"))
    out = _finetune(tmp_path, free_port(), ["--train_iters", "9", "--exit_signal_handler",
                                            "--fault_inject", "kill:0:3", "--save_interval", "100"])
    assert "exiting program after receiving SIGTERM" in out, out[-1500:]
    saved = open(tmp_path / "ckpt" / "latest_checkpointed_iteration.txt").read().strip()
    assert saved in ("3", "4"), saved
    assert not re.search(r"iteration\s+8/", out)
    out2 = _finetune(tmp_path, free_port(), ["--train_iters", "9", "--load", str(tmp_path / "ckpt"),
                                             "--exit_interval", "6", "--save_interval", "100"])
    assert "exiting program at iteration 6" in out2, out2[-1500:]
    assert open(tmp_path / "ckpt" / "latest_checkpointed_iteration.txt").read().strip() == "6"


def test_non_finite_gradient_skips_the_update(tmp_path):
    """A NaN planted in one gradient (--fault_inject nan:<rank>:<iteration>): the step's gradient norm
    is reported as nan, the parameters are left untouched by that step, and training carries on with
    finite losses afterwards."""
    from tests.helpers.launch import free_port
    _corpus(str(tmp_path / "corpus"))
    out = _finetune(tmp_path, free_port(), ["--train_iters", "6", "--fault_inject", "nan:0:2", "--log_interval", "1"])
    losses = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+|nan)", out)]
    assert len(losses) == 6 and all(math.isfinite(x) for x in losses), out[-2500:]
    assert re.search(r"iteration\s+3/.*grad norm: nan", out), out[-2500:]


def test_pretrain_gpt_sample_based_with_batch_rampup(tmp_path):
    """pretrain_gpt.py (GPT-2 style: LayerNorm, GeLU, learned absolute positions, tied embeddings),
    sample-based schedule (--train_samples / --lr_decay_samples / --lr_warmup_samples) with a global
    batch size that ramps 2 -> 4 -> 6 (megatron/microbatches.py:78-144)."""
    from tests.helpers.launch import free_port
    _corpus(str(tmp_path / "corpus"))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    cmd = [sys.executable, os.path.join(ROOT, "pretrain_gpt.py"),
           "--num_layers", "2", "--hidden_size", "64", "--num_attention_heads", "4", "--seq_length", "32",
           "--max_position_embeddings", "32", "--micro_batch_size", "2", "--global_batch_size", "6",
           "--rampup_batch_size", "2", "2", "24", "--train_samples", "60", "--lr_decay_samples", "60",
           "--lr_warmup_samples", "8", "--lr", "1e-3", "--min_lr", "1e-4", "--lr_decay_style", "linear",
           "--tokenizer_type", "NullTokenizer", "--vocab_size", "1024", "--data_path", str(tmp_path / "corpus"),
           "--split", "90,5,5", "--distributed_backend", "gloo", "--log_interval", "1", "--eval_interval", "1000",
           "--eval_iters", "1", "--num_workers", "0", "--use_cpu_initialization", "--no_bias_gelu_fusion",
           "--no_masked_softmax_fusion", "--hidden_dropout", "0", "--attention_dropout", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    gbs = [int(x) for x in re.findall(r"global batch size:\s+(\d+)", r.stdout)]
    consumed = [int(x) for x in re.findall(r"consumed samples:\s+(\d+)", r.stdout)]
    assert gbs[0] == 2 and gbs[-1] == 6 and sorted(set(gbs)) == [2, 4, 6] and gbs == sorted(gbs), gbs
    assert consumed[-1] >= 60 and all(b - a == g for a, b, g in zip(consumed, consumed[1:], gbs[1:])), (consumed, gbs)
    losses = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", r.stdout)]
    # uniformly random tokens: nothing to learn beyond log(vocab) ~ 6.9, but it must stay there
    assert len(losses) == len(gbs) and all(6.0 < x < 8.0 for x in losses), losses
