"""Checkpoint -> resume through the public training path: a run that saves
after two steps and a second process that loads and continues must reproduce the loss curve of an
uninterrupted run, for the replicated optimizer (model_optim_rng.pt) and for the distributed
optimizer on two data-parallel ranks (model_rng.pt + per-rank optim.pt)."""
import os

import pytest

from tests.helpers.launch import run_tiny

KW = dict(layers=2, num_microbatches=2, backend="gloo", dtype="fp32", hidden=128, heads=4, kv_heads=2, seq=64)
ENV = {"B200_DISABLE_KERNELS": "1", "CUDA_VISIBLE_DEVICES": ""}


def _close(a, b, rel=1e-5):
    assert len(a) == len(b) > 0
    for x, y in zip(a, b):
        assert abs(x - y) <= rel * abs(y), (a, b)


@pytest.mark.parametrize("nproc,dist_opt", [(1, False), (2, True)])
def test_save_then_resume_matches_uninterrupted_run(tmp_path, nproc, dist_opt):
    kw = dict(KW, dist_opt=dist_opt, nproc=nproc, env_extra=ENV)
    full = run_tiny(steps=4, **kw)["losses"]
    ckpt = str(tmp_path / "ckpt")
    first = run_tiny(steps=2, save=ckpt, **kw)["losses"]
    _close(first, full[:2])
    tracker = os.path.join(ckpt, "latest_checkpointed_iteration.txt")
    assert open(tracker).read().strip() == "2"
    rank_dir = os.path.join(ckpt, "iter_0000002", "mp_rank_00")
    files = sorted(os.listdir(rank_dir)) if os.path.isdir(rank_dir) else sorted(os.listdir(os.path.join(ckpt, "iter_0000002")))
    if dist_opt:
        assert any("model_rng.pt" in f for f in files) or any("mp_rank_00" in f for f in files), files
    else:
        assert "model_optim_rng.pt" in files, files
    second = run_tiny(steps=2, load=ckpt, **kw)["losses"]
    _close(second, full[2:])


def _gpt_cli(tmp_path, nproc, extra, tp=1, pp=1):
    """finetune.py on a tiny GPT (learned positions, biases, tied embeddings) -> stdout."""
    import subprocess
    import sys
    from tests.helpers.launch import ROOT, free_port
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), **ENV)
    argv = ["--model_name", "gpt", "--position_embedding_type", "absolute", "--num_layers", "4", "--hidden_size", "64",
            "--num_attention_heads", "4", "--seq_length", "32", "--max_position_embeddings", "32",
            "--micro_batch_size", "2", "--global_batch_size", "4", "--lr", "1e-3", "--min_lr", "1e-4",
            "--lr_decay_style", "cosine", "--lr_decay_iters", "10", "--train_iters", "3",
            "--tokenizer_type", "NullTokenizer", "--vocab_size", "1021", "--data_path", str(tmp_path / "corpus"),
            "--split", "90,5,5", "--distributed_backend", "gloo", "--log_interval", "1", "--eval_interval", "100",
            "--eval_iters", "2", "--num_workers", "0", "--use_cpu_initialization", "--no_bias_gelu_fusion",
            "--no_masked_softmax_fusion", "--tensor_model_parallel_size", str(tp),
            "--pipeline_model_parallel_size", str(pp)] + (["--sequence_parallel"] if tp > 1 else []) + extra
    if nproc == 1:
        env.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
        cmd = [sys.executable, os.path.join(ROOT, "finetune.py")]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
               "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"], os.path.join(ROOT, "finetune.py")]
    r = subprocess.run(cmd + argv, capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r.stdout


def _test_loss(out):
    import re
    m = re.search(r"end of training for test data \| lm loss value: ([0-9.E+-]+)", out)
    assert m, out[-2000:]
    return float(m.group(1))


def test_trained_gpt_checkpoint_reshards_to_tp2_pp2_and_evaluates_the_same(tmp_path):
    """The user workflow behind tools/checkpoint_util.py: train on one rank and save; reshard the
    iteration checkpoint to tp 2 x pp 2; load it with the training entry point on four ranks. The
    resharded checkpoint carries the integer iteration (nothing is re-trained) and the test-set loss
    equals the one-rank loss of the same weights."""
    import subprocess
    import sys
    from tests.helpers.launch import ROOT
    from tests.test_cli_e2e_cpu import _corpus
    _corpus(str(tmp_path / "corpus"))
    trained = _gpt_cli(tmp_path, 1, ["--save", str(tmp_path / "ck"), "--save_interval", "3"])
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "checkpoint_util.py"), "--model_type", "GPT",
                        "--load_dir", str(tmp_path / "ck"), "--save_dir", str(tmp_path / "ck22"),
                        "--target_tensor_parallel_size", "2", "--target_pipeline_parallel_size", "2"],
                       capture_output=True, text=True, cwd=ROOT, env=dict(os.environ, **ENV))
    assert r.returncode == 0, r.stderr[-3000:]
    assert open(tmp_path / "ck22" / "latest_checkpointed_iteration.txt").read().strip() == "3"
    assert sorted(os.listdir(tmp_path / "ck22" / "iter_0000003")) == [
        "mp_rank_00_000", "mp_rank_00_001", "mp_rank_01_000", "mp_rank_01_001"]
    sharded = _gpt_cli(tmp_path, 4, ["--load", str(tmp_path / "ck22"), "--no_load_optim", "--no_load_rng"], tp=2, pp=2)
    assert "at iteration 3" in sharded and not [l for l in sharded.splitlines() if l.startswith(" iteration ")]
    assert abs(_test_loss(sharded) - _test_loss(trained)) <= 2e-5 * _test_loss(trained)


def test_gpt_tied_embeddings_on_four_pipeline_stages(tmp_path):
    """Tied input/output embeddings with MIDDLE pipeline stages present: only the first and last stage
    are in the embedding group (the middle ones must stay out of its all-reduces), and after training
    the last stage's head copy still equals the first stage's embedding table."""
    import torch
    from tests.test_cli_e2e_cpu import _corpus
    _corpus(str(tmp_path / "corpus"))
    out = _gpt_cli(tmp_path, 4, ["--save", str(tmp_path / "ck"), "--save_interval", "3", "--hidden_dropout", "0",
                                 "--attention_dropout", "0"], pp=4)
    import re
    losses = [float(x) for x in re.findall(r"lm loss: ([0-9.E+-]+)", out)]
    assert len(losses) >= 3 and losses[2] < losses[0], losses
    import megatron  # noqa: F401  (lets torch.load resolve the pickled argument classes)
    first = torch.load(tmp_path / "ck" / "iter_0000003" / "mp_rank_00_000" / "model_optim_rng.pt",
                       map_location="cpu", weights_only=False)["model"]
    last = torch.load(tmp_path / "ck" / "iter_0000003" / "mp_rank_00_003" / "model_optim_rng.pt",
                      map_location="cpu", weights_only=False)["model"]
    table = first["language_model"]["embedding"]["word_embeddings"]["weight"]
    head = last["word_embeddings_for_head"]["weight"]
    assert torch.equal(table, head)
    middle = torch.load(tmp_path / "ck" / "iter_0000003" / "mp_rank_00_001" / "model_optim_rng.pt",
                        map_location="cpu", weights_only=False)["model"]
    assert "word_embeddings_for_head" not in middle and "embedding" not in middle["language_model"]


def test_interleaved_pipeline_checkpoint_unshards_to_one_rank(tmp_path):
    """A checkpoint written by the interleaved schedule (two stages x two model chunks, tied embeddings)
    goes back to tp 1 x pp 1 through tools/checkpoint_util.py with the layers in their global order:
    the one-rank test loss equals the one the pipeline run reported for the same weights."""
    import subprocess
    import sys
    from tests.helpers.launch import ROOT
    from tests.test_cli_e2e_cpu import _corpus
    _corpus(str(tmp_path / "corpus"))
    trained = _gpt_cli(tmp_path, 2, ["--save", str(tmp_path / "ck"), "--save_interval", "3",
                                     "--num_layers_per_virtual_pipeline_stage", "1"], pp=2)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "checkpoint_util.py"), "--model_type", "GPT",
                        "--load_dir", str(tmp_path / "ck"), "--save_dir", str(tmp_path / "ck11"),
                        "--target_tensor_parallel_size", "1", "--target_pipeline_parallel_size", "1"],
                       capture_output=True, text=True, cwd=ROOT, env=dict(os.environ, **ENV))
    assert r.returncode == 0, r.stderr[-3000:]
    single = _gpt_cli(tmp_path, 1, ["--load", str(tmp_path / "ck11"), "--no_load_optim", "--no_load_rng"])
    assert "at iteration 3" in single
    assert abs(_test_loss(single) - _test_loss(trained)) <= 2e-5 * _test_loss(trained)
