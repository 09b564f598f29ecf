"""Downstream tasks on CPU: MNLI finetuning through tasks/main.py on a synthetic TSV, the
zero-shot LM windowing, answer matching and F1."""
import json
import math
import os
import subprocess
import sys

import numpy as np
import pytest

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


def _zeroshot(tmp_path, task, data, extra=()):
    """tasks/main.py --task LAMBADA | WIKITEXT103 with a random tiny GPT on CPU."""
    import random as _random
    from tests.helpers.launch import free_port
    words = ["alpha", "beta", "gamma", "delta", "kernel", "tensor", "memory", "stream", "token", "layer"]
    chars = sorted(set("".join(words) + " .=\n"))
    toks = [c if c != " " else "Ġ" for c in chars if c != "\n"] + ["Ċ", "<|endoftext|>"]
    (tmp_path / "vocab.json").write_text(json.dumps({t: i for i, t in enumerate(toks)}))
    (tmp_path / "merges.txt").write_text("#version: 0.2\n")
    rng = _random.Random(0)
    data(tmp_path, rng, words)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    cmd = [sys.executable, os.path.join(ROOT, "tasks", "main.py"), "--task", task,
           "--valid_data", str(tmp_path / "eval_data"), "--tokenizer_type", "GPT2BPETokenizer",
           "--vocab_file", str(tmp_path / "vocab.json"), "--merge_file", str(tmp_path / "merges.txt"),
           "--num_layers", "2", "--hidden_size", "64", "--num_attention_heads", "4", "--seq_length", "64",
           "--max_position_embeddings", "64", "--micro_batch_size", "4", "--log_interval", "100",
           "--distributed_backend", "gloo", "--use_cpu_initialization", "--no_load_optim", "--no_load_rng",
           "--num_workers", "0", *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-2500:] + r.stderr[-2500:]
    return r.stdout


def test_zeroshot_wikitext_perplexity_cli(tmp_path):
    def data(d, rng, words):
        lines = [" = Heading = \n"] + [" ".join(rng.choice(words) for _ in range(40)) + " . \n" for _ in range(30)]
        (d / "eval_data").write_text("".join(lines))
    out = _zeroshot(tmp_path, "WIKITEXT103", data, ["--overlapping_eval", "32"])
    import re
    m = re.search(r"avg loss: ([0-9.E+-]+) \| ppl: ([0-9.E+-]+) \| adjusted ppl: ([0-9.E+-]+) \| token ratio: ([0-9.]+)", out)
    assert m, out[-1500:]
    loss, ppl = float(m.group(1)), float(m.group(2))
    # an untrained model over a ~25-symbol vocabulary: loss near log(vocab), far from degenerate
    assert 1.5 < loss < 6.0 and abs(ppl - math.exp(loss)) / ppl < 1e-2 and float(m.group(4)) > 1.0


def test_zeroshot_lambada_accuracy_cli(tmp_path):
    def data(d, rng, words):
        rows = [{"text": " ".join(rng.choice(words) for _ in range(12))} for _ in range(10)]
        (d / "eval_data").write_text("\n".join(json.dumps(r) for r in rows) + "\n")
    out = _zeroshot(tmp_path, "LAMBADA", data)
    import re
    m = re.search(r"number correct: ([0-9.E+-]+) \| total examples: ([0-9.E+-]+) \| avg accuracy: ([0-9.E+-]+)", out)
    assert m, out[-1500:]
    assert float(m.group(2)) == 10 and 0.0 <= float(m.group(3)) <= 1.0


def test_race_finetune_cpu(tmp_path):
    """RACE through tasks/main.py: multiple-choice head over NUM_CHOICES (context, question + option)
    pairs per sample, one epoch of finetuning and the accuracy report."""
    words = "the of and to in a is that for it as was with be by on not he this are".split()
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words
    (tmp_path / "vocab.txt").write_text("\n".join(vocab) + "\n")
    rng = np.random.RandomState(1)

    def split(name, n_articles):
        d = tmp_path / "RACE" / name / "middle"
        d.mkdir(parents=True)
        for a in range(n_articles):
            doc = {"article": " ".join(rng.choice(words, 30)),
                   "questions": [" ".join(rng.choice(words, 5)) + " ?", "it was _ by the"],
                   "options": [[" ".join(rng.choice(words, 2)) for _ in range(4)] for _ in range(2)],
                   "answers": ["ABCD"[int(rng.randint(4))] for _ in range(2)]}
            (d / f"{a}.txt").write_text(json.dumps(doc) + "\n")
        return str(d)
    train, dev = split("train", 6), split("dev", 2)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1",
               B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, os.path.join(ROOT, "tasks", "main.py"), "--task", "RACE", "--epochs", "1",
           "--train_data", train, "--valid_data", dev,
           "--tokenizer_type", "BertWordPieceLowerCase", "--vocab_file", str(tmp_path / "vocab.txt"),
           "--num_layers", "2", "--hidden_size", "32", "--num_attention_heads", "2", "--seq_length", "64",
           "--max_position_embeddings", "64", "--micro_batch_size", "2", "--global_batch_size", "2",
           "--lr", "1e-3", "--lr_decay_style", "linear", "--lr_warmup_fraction", "0.1",
           "--distributed_backend", "gloo", "--use_cpu_initialization", "--no_bias_gelu_fusion",
           "--no_masked_softmax_fusion", "--num_workers", "0", "--log_interval", "2", "--eval_interval", "100",
           "--eval_iters", "1", "--make_vocab_size_divisible_by", "8", "--hidden_dropout", "0.0",
           "--attention_dropout", "0.0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "total number of samples: 12" in r.stdout and "total number of samples: 4" in r.stdout
    assert "overall: correct / total = " in r.stdout and "done :-)" in r.stdout


def test_msdp_preprocessing_and_f1(tmp_path):
    """Wizard-of-Wikipedia dump -> tab-separated turns + reference files, response-stage input assembly,
    prompt sampling, and the token-level F1 used to score both generation stages."""
    wow = [{"chosen_topic": "Kernels", "dialog": [
        {"speaker": "0_Apprentice", "text": "Tell me about\nGPU kernels."},
        {"speaker": "1_Wizard", "text": "A kernel runs on many threads.",
         "checked_sentence": {"s0": "A kernel is a function launched on the GPU."}, "checked_passage": {"p0": "CUDA"}},
        {"speaker": "0_Apprentice", "text": "How many threads?"},
        {"speaker": "1_Wizard", "text": "Thousands at once.", "checked_sentence": {}, "checked_passage": {}}]},
        {"chosen_topic": "Memory", "dialog": [
            {"speaker": "0_Wizard", "text": "Let us talk about memory."},
            {"speaker": "1_Apprentice", "text": "Is HBM fast?"},
            {"speaker": "0_Wizard", "text": "Yes, HBM3e delivers high bandwidth.",
             "checked_sentence": {"s": "HBM3e delivers high bandwidth."}, "checked_passage": {}}]}]
    raw = tmp_path / "wow.json"
    raw.write_text(json.dumps(wow))
    out, kref, rref = tmp_path / "proc.txt", tmp_path / "knwl_ref.txt", tmp_path / "resp_ref.txt"
    pre = os.path.join(ROOT, "tasks", "msdp", "preprocessing.py")

    def run(*a):
        r = subprocess.run([sys.executable, pre, *map(str, a)], capture_output=True, text=True, cwd=ROOT, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
    run("--func", "process_wow_dataset", "--raw_file", raw, "--processed_file", out, "--knwl_ref_file", kref,
        "--resp_ref_file", rref)
    rows = [l.rstrip("\n").split("\t") for l in open(out)]
    assert rows == [["CUDA", "Tell me about GPU kernels.", "A kernel is a function launched on the GPU.",
                     "A kernel runs on many threads."],
                    ["Memory", "Let us talk about memory. [SEP] Is HBM fast?", "HBM3e delivers high bandwidth.",
                     "Yes, HBM3e delivers high bandwidth."]]
    assert open(kref).read().splitlines() == [r[2] for r in rows]
    assert open(rref).read().splitlines() == [r[3] for r in rows]
    gen = tmp_path / "knwl_gen.txt"
    gen.write_text("A kernel is a GPU function.<|endoftext|>\nHBM is fast memory.\n")
    resp_in = tmp_path / "resp_in.txt"
    run("--func", "prepare_input", "--test_file", out, "--knwl_gen_file", gen, "--processed_file", resp_in)
    got = [l.rstrip("\n").split("\t") for l in open(resp_in)]
    assert [g[2] for g in got] == ["A kernel is a GPU function.", "HBM is fast memory."] and got[0][0] == "CUDA"
    prompts = tmp_path / "resp_prompts.txt"
    run("--func", "get_resp_gen_prompts", "--train_file", out, "--processed_file", prompts, "--seed", "7")
    # only turns whose response re-uses most of the knowledge sentence qualify as prompts
    assert open(prompts).read().splitlines() == [
        "Topic: Memory. User says: Is HBM fast? We know that: HBM3e delivers high bandwidth. "
        "System replies: Yes, HBM3e delivers high bandwidth."]
    from msdp.metrics import F1Metric
    p, r, f1 = F1Metric.compute_all_pairs(["The kernel is a GPU function!", "nothing in common"],
                                          ["a kernel is a function launched on the gpu", "HBM3e delivers"])
    # first pair: articles / punctuation dropped -> guess {kernel,is,gpu,function}, 4 of 4 in the answer
    assert 0.4 < p < 0.6 and 0.25 < r < 0.45 and 0.3 < f1 < 0.5
    assert F1Metric.compute_each_pair("same words here", "same words here")[2] == 1.0


@pytest.mark.parametrize("nproc", [1, 2])
def test_orqa_supervised_retriever_finetune_cli(tmp_path, nproc):
    """tasks/main.py --task RET-FINETUNE-NQ on DPR-format json: biencoder with in-batch negatives gathered
    over the data-parallel group plus one hard negative per query. With near-identical random embeddings
    the loss starts at log(#contexts scored per query) = log(nproc * (4 positives + 4 negatives)); the
    end-of-epoch callback reports the average rank and top-k accuracy on the validation file."""
    import math
    import re
    words = "the of and to in a is that for it as was with be by on not he this are".split()
    (tmp_path / "vocab.txt").write_text("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words) + "\n")
    rng = np.random.RandomState(0)

    def ctx():
        return {"title": " ".join(rng.choice(words, 2)), "text": " ".join(rng.choice(words, 8))}

    def rows(n):
        return [{"question": " ".join(rng.choice(words, 5)) + "?", "answers": [str(rng.choice(words))],
                 "positive_ctxs": [ctx()], "hard_negative_ctxs": [ctx(), ctx()], "negative_ctxs": [ctx(), ctx()]}
                for _ in range(n)]
    (tmp_path / "train.json").write_text(json.dumps(rows(16)))
    (tmp_path / "dev.json").write_text(json.dumps(rows(8)))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), B200_DISABLE_KERNELS="1",
               CUDA_VISIBLE_DEVICES="")
    if nproc == 1:
        env.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
        launcher = [sys.executable]
    else:
        launcher = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                    "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"]]
    cmd = launcher + [
        os.path.join(ROOT, "tasks", "main.py"), "--task", "RET-FINETUNE-NQ", "--epochs", "1",
        "--train_data", str(tmp_path / "train.json"), "--valid_data", str(tmp_path / "dev.json"),
        "--tokenizer_type", "BertWordPieceLowerCase", "--vocab_file", str(tmp_path / "vocab.txt"),
        "--num_layers", "2", "--hidden_size", "32", "--num_attention_heads", "2", "--seq_length", "32",
        "--retriever_seq_length", "32", "--max_position_embeddings", "32", "--micro_batch_size", "4",
        "--global_batch_size", str(4 * nproc), "--lr", "1e-3", "--lr_decay_style", "linear",
        "--lr_warmup_fraction", "0.1", "--distributed_backend", "gloo", "--use_cpu_initialization",
        "--no_bias_gelu_fusion", "--no_masked_softmax_fusion", "--num_workers", "0", "--log_interval", "1",
        "--eval_interval", "100", "--eval_iters", "1", "--make_vocab_size_divisible_by", "8",
        "--hidden_dropout", "0.0", "--attention_dropout", "0.0", "--retriever_report_topk_accuracies", "1", "2",
        "--retriever_score_scaling", "--train_with_neg", "--train_hard_neg", "1", "--val_av_rank_hard_neg", "2",
        "--val_av_rank_other_neg", "2", "--eval_micro_batch_size", "4", "--biencoder_projection_dim", "16"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    losses = [float(x) for x in re.findall(r"\| loss: ([0-9.E+-]+)", r.stdout)]
    assert len(losses) == 16 // (4 * nproc)
    assert abs(losses[0] - math.log(8 * nproc)) < 5e-3, losses
    assert re.search(r"epoch:0 name:dev rank: [0-9.]+ top1_acc: [0-9.]+ top2_acc: [0-9.]+", r.stdout), r.stdout[-800:]
    assert "done :-)" in r.stdout


def test_open_retrieval_pipeline_ict_pretrain_index_evaluate(tmp_path):
    """The open-retrieval workflow end to end: pretrain_ict.py saves a biencoder; IndexBuilder embeds an
    evidence TSV with the context tower on TWO data-parallel ranks (each dumps a shard, rank 0 folds
    the other's into its own and writes --embedding_path); tasks/main.py --task RETRIEVER-EVAL embeds
    the questions with the query tower, searches the index and reports top-k answer hits."""
    import pickle
    import re
    sys.path.insert(0, os.path.join(ROOT, "tests", "helpers"))
    try:
        import legacy_worker
    finally:
        sys.path.pop(0)
    vf, prefix, titles = legacy_worker.make_corpus(str(tmp_path))
    words = [w for w in open(vf).read().split() if not w.startswith("[")][:30]
    rng = np.random.RandomState(0)
    model = (f"--num_layers 2 --hidden_size 64 --num_attention_heads 4 --seq_length 48 --max_position_embeddings 64 "
             f"--micro_batch_size 4 --tokenizer_type BertWordPieceLowerCase --vocab_file {vf} "
             f"--distributed_backend gloo --use_cpu_initialization --hidden_dropout 0.0 --attention_dropout 0.0 "
             f"--no_bias_gelu_fusion --no_masked_softmax_fusion --num_workers 0 --make_vocab_size_divisible_by 8 "
             f"--biencoder_projection_dim 32 --retriever_score_scaling --retriever_report_topk_accuracies 1 2 5 "
             f"--retriever_seq_length 48").split()

    def run(script, extra, nproc=1):
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), B200_DISABLE_KERNELS="1",
                   CUDA_VISIBLE_DEVICES="")
        if nproc == 1:
            env.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
            launcher = [sys.executable]
        else:
            launcher = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                        "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"]]
        r = subprocess.run(launcher + [script] + model + extra.split(), capture_output=True, text=True,
                           timeout=900, env=env, cwd=ROOT)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3500:]
        return r.stdout

    ckpt = tmp_path / "ckpt"
    run(os.path.join(ROOT, "pretrain_ict.py"),
        f"--global_batch_size 4 --train_iters 4 --lr 2e-3 --min_lr 1e-4 --lr_decay_style constant --data_path {prefix} "
        f"--titles_data_path {titles} --split 90,5,5 --eval_iters 1 --eval_interval 2 --log_interval 2 "
        f"--data_impl mmap --seed 7 --save {ckpt} --save_interval 4 --query_in_block_prob 0.2 --ict_head_size 32")
    with open(tmp_path / "evidence.tsv", "w") as f:
        f.write("id\ttext\ttitle\n")
        for i in range(1, 21):
            f.write(f"{i}\t{' '.join(rng.choice(words, 12))}\t{' '.join(rng.choice(words, 2))}\n")
    with open(tmp_path / "qa.csv", "w") as f:
        for _ in range(6):
            f.write(f"{' '.join(rng.choice(words, 5))}?\t{json.dumps([str(rng.choice(words))])}\n")
    (tmp_path / "build_index.py").write_text(
        f"import sys\nsys.path.insert(0, {ROOT!r})\n"
        "from epfl_megatron_b200.initialize import initialize_megatron\n"
        "from epfl_megatron_b200.indexer import IndexBuilder\n"
        "initialize_megatron()\nIndexBuilder().build_and_save_index()\n")
    shared = (f"--load {ckpt} --ict_load {ckpt} --evidence_data_path {tmp_path / 'evidence.tsv'} "
              f"--embedding_path {tmp_path / 'emb.pkl'}")
    out = run(str(tmp_path / "build_index.py"), shared + " --indexer_batch_size 4 --indexer_log_interval 1", nproc=2)
    assert "Finished merging 2 shards for a total of 20 embeds" in out, out[-1500:]
    with open(tmp_path / "emb.pkl", "rb") as f:
        embeds = pickle.load(f)["embed_data"]
    assert sorted(embeds) == list(range(1, 21)) and all(v.shape == (32,) for v in embeds.values())
    assert not os.path.exists(tmp_path / "emb_tmp")
    out = run(os.path.join(ROOT, "tasks", "main.py"),
              "--task RETRIEVER-EVAL " + shared + f" --qa_data_dev {tmp_path / 'qa.csv'} --faiss_topk_retrievals 5 "
              "--faiss_match string")
    hits = re.search(r"topk-5 documents hits \[([0-9, ]+)\]", out)
    assert hits, out[-1500:]
    hits = [int(x) for x in hits.group(1).split(",")]
    assert len(hits) == 5 and hits == sorted(hits) and hits[-1] <= 6
    assert re.search(r"top-5: [0-9.]+", out) and "done :-)" in out


@pytest.mark.parametrize("prompt_type", ["knowledge", "response"])
def test_msdp_prompt_cli_generates_one_line_per_sample(tmp_path, prompt_type):
    """tasks/msdp/main.py --task MSDP-PROMPT with a locally built (random, tiny) GPT: the few-shot prompt
    is assembled per test sample, greedy generation runs through the public generation API and the
    first generated line of every sample lands in --sample_output_file."""
    words = ["alpha", "beta", "gamma", "delta", "kernel", "tensor", "memory", "stream", "token", "layer"]
    text = "".join(words) + " .=()>:Topic: . User says: We know that: System replies:"
    toks = list(dict.fromkeys([c if c != " " else "Ġ" for c in sorted(set(text))] + ["Ċ", "<|endoftext|>"]))
    (tmp_path / "vocab.json").write_text(json.dumps({t: i for i, t in enumerate(toks)}))
    (tmp_path / "merges.txt").write_text("#version: 0.2\n")
    samples = [("alpha", "beta gamma", "delta kernel"), ("tensor", "memory stream", "token layer"),
               ("layer", "token alpha", "beta")]
    with open(tmp_path / "test.txt", "w") as f:
        for topic, last, know in samples:
            f.write(f"{topic}\talpha [SEP] {last}\t{know}\tgamma\n")
    if prompt_type == "knowledge":
        with open(tmp_path / "prompts", "w") as f:
            for topic, last, know in samples:
                f.write(json.dumps({f"{topic} {last}": [f"( {last} ) {topic} => {know}"] * 2}) + "\n")
    else:
        (tmp_path / "prompts").write_text(
            "Topic: alpha. User says: beta We know that: gamma System replies: delta\n" * 2)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    cmd = [sys.executable, os.path.join(ROOT, "tasks", "msdp", "main.py"), "--task", "MSDP-PROMPT",
           "--prompt_type", prompt_type, "--prompt_file", str(tmp_path / "prompts"),
           "--sample_input_file", str(tmp_path / "test.txt"), "--sample_output_file", str(tmp_path / "out.txt"),
           "--num_prompt_examples", "2", "--out_seq_length", "8", "--tokenizer_type", "GPT2BPETokenizer",
           "--vocab_file", str(tmp_path / "vocab.json"), "--merge_file", str(tmp_path / "merges.txt"),
           "--num_layers", "2", "--hidden_size", "64", "--num_attention_heads", "4", "--seq_length", "256",
           "--max_position_embeddings", "256", "--micro_batch_size", "1", "--distributed_backend", "gloo",
           "--use_cpu_initialization", "--no_load_optim", "--no_load_rng", "--num_workers", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    lines = open(tmp_path / "out.txt").read().split("\n")
    assert len(lines) == len(samples) + 1 and lines[-1] == ""
    assert all(len(ln) <= 8 for ln in lines)        # at most --out_seq_length single-character tokens


def test_qqp_finetune_then_evaluation_only(tmp_path):
    """tasks/main.py --task QQP: one finetuning epoch with a checkpoint at the end, then an evaluation-only
    run (--epochs 0 --pretrained_checkpoint) of the saved weights: same accuracy line, predictions written
    next to the checkpoint."""
    import re
    import torch
    words = "the of and to in a is that for it as was with be by on not he this are".split()
    (tmp_path / "vocab.txt").write_text("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words) + "\n")
    rng = np.random.RandomState(0)

    def tsv(path, n):
        rows = ["id\tqid1\tqid2\tquestion1\tquestion2\tis_duplicate"]
        for i in range(n):
            rows.append("\t".join([str(i), str(2 * i), str(2 * i + 1), " ".join(rng.choice(words, 6)),
                                   " ".join(rng.choice(words, 5)), str(i % 2)]))
        path.write_text("\n".join(rows) + "\n")
    tsv(tmp_path / "train.tsv", 24)
    tsv(tmp_path / "dev.tsv", 8)

    def run(extra):
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1",
                   B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
        cmd = [sys.executable, os.path.join(ROOT, "tasks", "main.py"), "--task", "QQP",
               "--train_data", str(tmp_path / "train.tsv"), "--valid_data", str(tmp_path / "dev.tsv"),
               "--tokenizer_type", "BertWordPieceLowerCase", "--vocab_file", str(tmp_path / "vocab.txt"),
               "--num_layers", "2", "--hidden_size", "32", "--num_attention_heads", "2", "--seq_length", "32",
               "--max_position_embeddings", "32", "--micro_batch_size", "4", "--global_batch_size", "4",
               "--lr", "1e-3", "--lr_decay_style", "linear", "--lr_warmup_fraction", "0.1",
               "--distributed_backend", "gloo", "--use_cpu_initialization", "--no_bias_gelu_fusion",
               "--no_masked_softmax_fusion", "--num_workers", "0", "--log_interval", "2", "--eval_interval", "100",
               "--eval_iters", "1", "--make_vocab_size_divisible_by", "8", "--hidden_dropout", "0.0",
               "--attention_dropout", "0.0"] + extra
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
        assert "done :-)" in r.stdout
        return re.search(r"overall: correct / total = (\d+) / (\d+)", r.stdout).groups()

    trained = run(["--epochs", "1", "--save", str(tmp_path / "ck"), "--save_interval", "100"])
    assert trained[1] == "8"
    assert open(tmp_path / "ck" / "latest_checkpointed_iteration.txt").read().strip() == "6"
    evaluated = run(["--epochs", "0", "--pretrained_checkpoint", str(tmp_path / "ck")])
    assert evaluated == trained
    preds = torch.load(tmp_path / "ck" / "predictions.pt", weights_only=False)
    (name, (softmaxes, labels, ids)), = preds.items()
    assert len(labels) == 8 and len(ids) == 8 and np.asarray(softmaxes).shape == (8, 2)
