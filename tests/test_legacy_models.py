"""BERT / T5 / ICT-biencoder models, their datasets and entry-point functions: a few CPU
training steps each on a toy WordPiece corpus (the reference has no automated test for these)."""
import json
import os
import subprocess
import sys

import pytest

from tests.helpers.launch import free_port

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(mode, tmp_path):
    env = dict(os.environ, MASTER_PORT=str(free_port()), B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "helpers", "legacy_worker.py"), mode,
                        str(tmp_path)], capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-5000:]
    return json.loads(lines[-1][7:])


@pytest.mark.parametrize("mode,keys", [
    ("bert", ["binary_head", "language_model", "lm_head"]),
    ("t5", ["language_model", "lm_head"]),
    ("ict", ["context_model", "query_model"]),
])
def test_legacy_model_trains(mode, keys, tmp_path):
    out = _run(mode, tmp_path)
    assert out["keys"] == keys
    losses = out["losses"]
    assert all(x == x and x < 50 for x in losses)
    assert min(losses[3:]) < losses[0], losses


@pytest.mark.parametrize("script,extra", [
    ("pretrain_bert.py", ""),
    ("pretrain_t5.py", "--decoder_seq_length 32 --vocab_extra_ids 100 --kv_channels 16 --ffn_hidden_size 128"),
    ("pretrain_ict.py", "--titles_data_path {titles} --query_in_block_prob 0.2 --ict_head_size 32 "
                        "--biencoder_projection_dim 32 --retriever_report_topk_accuracies 1 2 --retriever_score_scaling"),
])
def test_legacy_entry_points_from_the_command_line(script, extra, tmp_path):
    """The three legacy entry points as a user starts them: pretrain() with evaluation, logging and a
    checkpoint at the end."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "helpers"))
    try:
        import legacy_worker
    finally:
        sys.path.pop(0)
    vf, prefix, titles = legacy_worker.make_corpus(str(tmp_path))
    argv = (f"--num_layers 2 --hidden_size 64 --num_attention_heads 4 --seq_length 48 --max_position_embeddings 64 "
            f"--micro_batch_size 4 --global_batch_size 4 --train_iters 4 --lr 2e-3 --min_lr 1e-4 "
            f"--lr_decay_style constant --tokenizer_type BertWordPieceLowerCase --vocab_file {vf} --data_path {prefix} "
            f"--split 90,5,5 --distributed_backend gloo --use_cpu_initialization --hidden_dropout 0.0 "
            f"--attention_dropout 0.0 --no_bias_gelu_fusion --no_masked_softmax_fusion --num_workers 0 "
            f"--eval_iters 1 --eval_interval 2 --log_interval 2 --make_vocab_size_divisible_by 8 --data_impl mmap "
            f"--seed 7 --save {tmp_path / 'ckpt'} --save_interval 4 " + extra.format(titles=titles)).split()
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1",
               LOCAL_RANK="0", B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, script)] + argv, capture_output=True, text=True,
                       timeout=900, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2500:] + r.stderr[-3500:]
    assert "iteration        4/       4" in r.stdout and "validation loss at iteration 2" in r.stdout, r.stdout[-2000:]
    assert open(tmp_path / "ckpt" / "latest_checkpointed_iteration.txt").read().strip() == "4"


@pytest.mark.parametrize("stages,split,layers", [(2, 1, 2), (4, 2, 4)])
def test_t5_encoder_decoder_pipeline_split(tmp_path, stages, split, layers):
    """pretrain_t5.py with the encoder on the pipeline ranks before --pipeline_model_parallel_split_rank
    and the decoder after it. --decoder_num_layers is not given (the reference's T5 recipes pass only
    --num_layers): the decoder is as deep as the encoder. With four stages the middle ranks hold
    neither embedding copy and stay out of the embedding-group all-reduces."""
    import re
    sys.path.insert(0, os.path.join(ROOT, "tests", "helpers"))
    try:
        import legacy_worker
    finally:
        sys.path.pop(0)
    vf, prefix, _ = legacy_worker.make_corpus(str(tmp_path))
    argv = (f"--num_layers {layers} --hidden_size 64 --num_attention_heads 4 --seq_length 48 --max_position_embeddings 64 "
            f"--micro_batch_size 2 --global_batch_size 4 --train_iters 3 --lr 2e-3 --min_lr 1e-4 "
            f"--lr_decay_style constant --tokenizer_type BertWordPieceLowerCase --vocab_file {vf} --data_path {prefix} "
            f"--split 90,5,5 --distributed_backend gloo --use_cpu_initialization --hidden_dropout 0.0 "
            f"--attention_dropout 0.0 --no_bias_gelu_fusion --no_masked_softmax_fusion --num_workers 0 "
            f"--eval_iters 1 --eval_interval 2 --log_interval 1 --make_vocab_size_divisible_by 8 --data_impl mmap "
            f"--seed 7 --decoder_seq_length 32 --vocab_extra_ids 100 --kv_channels 16 --ffn_hidden_size 128 "
            f"--pipeline_model_parallel_size {stages} --pipeline_model_parallel_split_rank {split}").split()
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), B200_DISABLE_KERNELS="1",
               CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={stages}",
           "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"], os.path.join(ROOT, "pretrain_t5.py")]
    r = subprocess.run(cmd + argv, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3500:]
    losses = [float(x) for x in re.findall(r"\| lm loss: ([0-9.E+-]+)", r.stdout)]
    assert len(losses) >= 3 and losses[-1] < losses[0], r.stdout[-1500:]
    assert "validation loss at iteration 2" in r.stdout


def test_bert_checkpoint_reshards_with_its_heads(tmp_path):
    """tools/checkpoint_util.py --model_type BERT carries the masked-LM head (its bias is vocabulary
    parallel), the binary head and the pooler: continuing from the TP2 reshard on two ranks starts from
    the same losses as continuing from the original on one rank. A BERT checkpoint without its heads is
    refused instead of silently training re-initialised ones."""
    import re
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests", "helpers"))
    try:
        import legacy_worker
    finally:
        sys.path.pop(0)
    vf, prefix, _ = legacy_worker.make_corpus(str(tmp_path))

    def run(nproc, extra, expect_ok=True):
        argv = (f"--num_layers 2 --hidden_size 64 --num_attention_heads 4 --seq_length 48 --max_position_embeddings 64 "
                f"--micro_batch_size 2 --global_batch_size 4 --lr 2e-3 --min_lr 1e-4 --lr_decay_style constant "
                f"--tokenizer_type BertWordPieceLowerCase --vocab_file {vf} --data_path {prefix} --split 90,5,5 "
                f"--distributed_backend gloo --use_cpu_initialization --hidden_dropout 0.0 --attention_dropout 0.0 "
                f"--no_bias_gelu_fusion --no_masked_softmax_fusion --num_workers 0 --eval_iters 1 --eval_interval 100 "
                f"--log_interval 1 --make_vocab_size_divisible_by 8 --data_impl mmap --seed 7 " + extra).split()
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), B200_DISABLE_KERNELS="1",
                   CUDA_VISIBLE_DEVICES="")
        if nproc == 1:
            env.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
            cmd = [sys.executable, os.path.join(ROOT, "pretrain_bert.py")]
        else:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                   "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"],
                   os.path.join(ROOT, "pretrain_bert.py")]
        r = subprocess.run(cmd + argv, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
        if not expect_ok:
            return r
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3500:]
        return [(float(a), float(b)) for a, b in
                re.findall(r"lm loss: ([0-9.E+-]+) \| sop loss: ([0-9.E+-]+)", r.stdout)]

    run(1, f"--train_iters 3 --save {tmp_path / 'ck'} --save_interval 3")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "checkpoint_util.py"), "--model_type", "BERT",
                        "--load_dir", str(tmp_path / "ck"), "--save_dir", str(tmp_path / "ck_tp2"),
                        "--target_tensor_parallel_size", "2", "--target_pipeline_parallel_size", "1"],
                       capture_output=True, text=True, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    one = run(1, f"--train_iters 2 --finetune --load {tmp_path / 'ck'}")
    two = run(2, f"--train_iters 2 --finetune --load {tmp_path / 'ck_tp2'} --tensor_model_parallel_size 2")
    assert len(one) == len(two) == 2
    for (a_lm, a_sop), (b_lm, b_sop) in zip(one, two):
        assert abs(a_lm - b_lm) <= 1e-4 * a_lm and abs(a_sop - b_sop) <= 1e-4 * abs(a_sop), (one, two)
    # the loader -> saver plug-in protocol writes the same shards, heads and pooler included
    import queue as pyqueue
    import types
    sys.path.insert(0, ROOT)
    from tools import checkpoint_loader_megatron as loader, checkpoint_saver_megatron as saver
    q = pyqueue.Queue()
    a = types.SimpleNamespace(model_type="BERT", load_dir=str(tmp_path / "ck"), save_dir=str(tmp_path / "ck_queue"),
                              checking=True, true_vocab_size=None, target_tensor_parallel_size=2,
                              target_pipeline_parallel_size=1)
    loader.load_checkpoint(q, a)
    saver.save_checkpoint(q, a)

    def flat(x, pre=""):
        out = {}
        for k, val in x.items():
            if isinstance(val, dict):
                out.update(flat(val, pre + k + "/"))
            elif torch.is_tensor(val):
                out[pre + k] = val
        return out
    for t in range(2):
        d = flat(torch.load(tmp_path / "ck_tp2" / "iter_0000003" / f"mp_rank_{t:02d}" / "model_optim_rng.pt",
                            weights_only=False)["model"])
        v = flat(torch.load(tmp_path / "ck_queue" / "iter_0000003" / f"mp_rank_{t:02d}" / "model_optim_rng.pt",
                            weights_only=False)["model"])
        assert d.keys() == v.keys() and {"lm_head/bias", "binary_head/weight", "language_model/pooler/dense.weight",
                                         "language_model/embedding/tokentype_embeddings/weight"} <= d.keys(), sorted(d)
        assert all(torch.equal(d[k], v[k]) for k in d)
    # strip the heads from a copy of the original: loading must fail loudly
    import megatron  # noqa: F401
    path = tmp_path / "ck" / "iter_0000003" / "mp_rank_00" / "model_optim_rng.pt"
    sd = torch.load(path, map_location="cpu", weights_only=False)
    del sd["model"]["lm_head"]
    torch.save(sd, path)
    bad = run(1, f"--train_iters 2 --finetune --load {tmp_path / 'ck'}", expect_ok=False)
    assert bad.returncode != 0 and "lm_head is missing from the checkpoint" in bad.stderr


@pytest.mark.parametrize("script,extra", [
    ("pretrain_bert.py", ""),
    ("pretrain_t5.py", "--decoder_seq_length 32 --vocab_extra_ids 100 --kv_channels 16 --ffn_hidden_size 128 "),
])
def test_legacy_resume_replays_dropout(tmp_path, script, extra):
    """Save after two of four iterations (--exit_interval), resume in a new process: with dropout ON the
    loss curve equals the uninterrupted run digit for digit -- generator states, data position and
    optimizer state all come back (the data-loader iterators do not disturb the restored generator)."""
    import re
    sys.path.insert(0, os.path.join(ROOT, "tests", "helpers"))
    try:
        import legacy_worker
    finally:
        sys.path.pop(0)
    vf, prefix, _ = legacy_worker.make_corpus(str(tmp_path))

    def run(more):
        argv = (f"--num_layers 2 --hidden_size 64 --num_attention_heads 4 --seq_length 48 --max_position_embeddings 64 "
                f"--micro_batch_size 2 --global_batch_size 4 --lr 2e-3 --min_lr 1e-4 --train_iters 4 "
                f"--lr_decay_style constant --tokenizer_type BertWordPieceLowerCase --vocab_file {vf} "
                f"--data_path {prefix} --split 90,5,5 --distributed_backend gloo --use_cpu_initialization "
                f"--hidden_dropout 0.1 --attention_dropout 0.1 --no_bias_gelu_fusion --no_masked_softmax_fusion "
                f"--num_workers 0 --eval_iters 1 --eval_interval 100 --log_interval 1 "
                f"--make_vocab_size_divisible_by 8 --data_impl mmap --seed 7 " + extra + more).split()
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1",
                   LOCAL_RANK="0", B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
        r = subprocess.run([sys.executable, os.path.join(ROOT, script)] + argv, capture_output=True, text=True,
                           timeout=900, env=env, cwd=ROOT)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3500:]
        return re.findall(r"lm loss: ([0-9.E+-]+)", r.stdout)

    whole = run("")
    first = run(f"--save {tmp_path / 'ck'} --save_interval 2 --exit_interval 2")
    second = run(f"--load {tmp_path / 'ck'}")
    assert len(whole) == 4 and first == whole[:2] and second == whole[2:], (whole, first, second)
