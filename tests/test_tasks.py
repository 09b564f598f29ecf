"""Downstream tasks on CPU: MNLI finetuning through tasks/main.py on a synthetic TSV, the
zero-shot LM windowing, answer matching and F1."""
import os
import subprocess
import sys

import numpy as np

from tests.helpers.launch import free_port

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tasks"))


def test_lm_dataset_windows_cover_every_token_once():
    from zeroshot_gpt.datasets import _LMDataset
    tokens = list(range(1, 54))
    ds = _LMDataset(tokens, seq_len=16, pad_idx=0, num_original_tokens=40, num_tokenized_tokens=len(tokens),
                    overlapping_eval=8)
    scored = np.zeros(len(tokens), dtype=np.int64)
    for i in range(len(ds)):
        s = ds[i]
        start = i * 8
        for j, m in enumerate(s["pad_mask"]):
            if m:
                scored[start + 1 + j] += 1
    assert scored[0] == 0 and (scored[1:] == 1).all()     # every target token scored exactly once


def test_answer_matching_and_f1():
    from msdp.metrics import F1Metric
    from orqa.unsupervised.qa_utils import calculate_matches
    docs = {1: ("Paris is the capital of France.", "France"), 2: ("Berlin is in Germany.", "Germany")}
    stats = calculate_matches(docs, [["Paris"], ["Rome"]], [([2, 1], [0.9, 0.8]), ([1, 2], [0.5, 0.4])],
                              workers_num=0, match_type="string")
    assert stats.top_k_hits == [0, 1]
    p, r, f = F1Metric.compute_all_pairs(["the cat sat"], ["a cat sat down"])
    assert abs(f - 0.8) < 1e-9


def test_mnli_finetune_cpu(tmp_path):
    words = "the of and to in a is that for it as was with be by on not he this are".split()
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words
    (tmp_path / "vocab.txt").write_text("\n".join(vocab) + "\n")
    rng = np.random.RandomState(0)
    labels = ["contradiction", "entailment", "neutral"]

    def tsv(path, n):
        rows = ["\t".join(f"c{i}" for i in range(12))]
        for i in range(n):
            a = " ".join(rng.choice(words, 6))
            b = " ".join(rng.choice(words, 5))
            rows.append("\t".join([str(i)] + ["x"] * 7 + [a, b, "x", labels[i % 3]]))
        path.write_text("\n".join(rows) + "\n")
    (tmp_path / "MNLI").mkdir()
    tsv(tmp_path / "MNLI" / "train.tsv", 24)
    tsv(tmp_path / "MNLI" / "dev_matched.tsv", 8)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1",
               B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, os.path.join(ROOT, "tasks", "main.py"), "--task", "MNLI", "--epochs", "1",
           "--train_data", str(tmp_path / "MNLI" / "train.tsv"), "--valid_data", str(tmp_path / "MNLI" / "dev_matched.tsv"),
           "--tokenizer_type", "BertWordPieceLowerCase", "--vocab_file", str(tmp_path / "vocab.txt"),
           "--num_layers", "2", "--hidden_size", "32", "--num_attention_heads", "2", "--seq_length", "32",
           "--max_position_embeddings", "32", "--micro_batch_size", "4", "--global_batch_size", "4",
           "--lr", "1e-3", "--lr_decay_style", "linear", "--lr_warmup_fraction", "0.1",
           "--distributed_backend", "gloo", "--use_cpu_initialization", "--no_bias_gelu_fusion",
           "--no_masked_softmax_fusion", "--num_workers", "0", "--log_interval", "2", "--eval_interval", "100",
           "--eval_iters", "1", "--make_vocab_size_divisible_by", "8", "--hidden_dropout", "0.0",
           "--attention_dropout", "0.0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "overall: correct / total = " in r.stdout and "done :-)" in r.stdout
