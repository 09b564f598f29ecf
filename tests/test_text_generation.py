"""Text generation on CPU: sampling filters, beam pool, and (in a subprocess that owns the
global args) KV-cache decoding vs un-cached recomputation, scoring, beam search, REST server."""
import json
import os
import subprocess
import sys

import torch

from epfl_megatron_b200.text_generation.beam_utils import BeamHypotheses
from epfl_megatron_b200.text_generation.sampling import (modify_logits_for_top_k_filtering,
                                                         modify_logits_for_top_p_filtering, sample)
from epfl_megatron_b200.text_generation_server import validate_request
from tests.helpers.launch import free_port

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_top_k_top_p_filters():
    logits = torch.tensor([[1.0, 3.0, 2.0, 0.0], [0.5, 0.1, 0.2, 4.0]])
    a = logits.clone()
    modify_logits_for_top_k_filtering(a, 2)
    assert torch.isinf(a[0, 0]) and torch.isinf(a[0, 3]) and a[0, 1] == 3.0 and a[0, 2] == 2.0
    b = logits.clone()
    modify_logits_for_top_p_filtering(b, 0.5)
    assert b[0, 1] == 3.0 and torch.isinf(b[0, 3])       # the top token always survives
    assert b[1, 3] == 4.0 and torch.isinf(b[1, 0])
    assert sample(logits, top_k=1).tolist() == [1, 3]
    torch.manual_seed(0)
    s = sample(logits, top_k=2, temperature=0.7)
    assert s[0] in (1, 2) and s[1] in (0, 3)
    assert int(sample(torch.tensor([[0.0, 0.0, 9.0]]), top_k=1, vocab_size=2)) == 1   # clamp to vocab


def test_beam_hypotheses():
    h = BeamHypotheses(2, length_penalty=1.0)
    h.add("a", -4.0, 4)
    h.add("b", -2.0, 4)
    h.add("c", -8.0, 4)          # worse than both: dropped
    assert sorted(x[1] for x in h.beams) == ["a", "b"]
    assert h.is_done(-6.0, 4) and not h.is_done(-1.0, 4)


def test_request_validation():
    assert validate_request({})[1] == "prompts argument required"
    assert validate_request({"prompts": ["x"], "top_k": 5, "top_p": 0.5})[1].startswith("cannot set both")
    assert validate_request({"prompts": [""], "tokens_to_generate": 2})[1] == "Empty prompts require add_BOS=true"
    assert validate_request({"prompts": ["a", "b"], "beam_width": 2})[1].startswith("When doing beam_search")
    p, err = validate_request({"prompts": ["a"], "tokens_to_generate": 5, "top_p": 0.9, "top_p_decay": 0.5})
    assert err is None and p["top_p_decay"] == 0.5 and p["tokens_to_generate"] == 5
    assert validate_request({"prompts": ["a"], "tokens_to_generate": 0})[1].startswith("tokens_to_generate=0")


def test_generation_end_to_end_cpu():
    env = dict(os.environ, MASTER_PORT=str(free_port()), GEN_TEST_PORT=str(free_port()),
               B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "helpers", "gen_worker.py")],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-4000:]
    out = json.loads(lines[-1][7:])
    assert out["shape"] == [2, 13]
    for k in ("greedy_matches_uncached", "logprobs_finite", "sampled_texts", "score_shape_ok", "beam_ok",
              "server_ok", "server_rejects", "server_beam"):
        assert out[k], (k, out)


def test_server_cli_tensor_parallel_and_client(tmp_path):
    """tools/run_text_generation_server.py under torchrun with two tensor-parallel ranks: rank 0 serves
    HTTP, rank 1 sits in the worker loop and is driven by the broadcast opcodes; queried through the
    interactive client tools/text_generation_cli.py (megatron/text_generation_server.py:22-29,
    tools/run_text_generation_server.py:49-84)."""
    import socket
    import time
    port = free_port()
    env = dict(os.environ, B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(free_port()), os.path.join(ROOT, "tools", "run_text_generation_server.py"),
           "--model_name", "llama2", "--tensor_model_parallel_size", "2", "--num_layers", "2", "--hidden_size", "64",
           "--num_attention_heads", "4", "--num_attention_heads_kv", "2", "--ffn_hidden_size", "128",
           "--seq_length", "64", "--max_position_embeddings", "64", "--micro_batch_size", "1", "--use_rms_norm",
           "--glu_activation", "swiglu", "--no_tie_embed_logits", "--position_embedding_type", "rotary",
           "--no_bias_gelu_fusion", "--tokenizer_type", "NullTokenizer", "--vocab_size", "512",
           "--distributed_backend", "gloo", "--use_cpu_initialization", "--host", "127.0.0.1", "--port", str(port)]
    log = open(tmp_path / "server.log", "w")
    server = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, env=env, cwd=ROOT)
    try:
        deadline = time.time() + 180
        while time.time() < deadline:
            assert server.poll() is None, open(tmp_path / "server.log").read()[-3000:]
            try:
                socket.create_connection(("127.0.0.1", port), timeout=1).close()
                break
            except OSError:
                time.sleep(1)
        else:
            raise AssertionError("server did not come up:\n" + open(tmp_path / "server.log").read()[-3000:])
        client = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "text_generation_cli.py"),
                                 f"127.0.0.1:{port}"], input="5 6 7 8\n6\n9 10\n3\n", capture_output=True, text=True,
                                timeout=300, cwd=ROOT)
        assert client.returncode == 0, client.stderr[-2000:]
        answers = [ln for ln in client.stdout.splitlines() if ln and ln[0].isdigit()]
        assert len(answers) == 2, client.stdout
        first, second = answers[0].split(), answers[1].split()
        assert first[:4] == ["5", "6", "7", "8"] and len(first) == 10 and second[:2] == ["9", "10"] and len(second) == 5
    finally:
        server.terminate()
        try:
            server.wait(timeout=30)
        except subprocess.TimeoutExpired:
            server.kill()
        log.close()


def _serve_and_query(tmp_path, tag, nproc, model_args, extra, bodies):
    """Start tools/run_text_generation_server.py on ``nproc`` ranks, PUT every body to /api, stop it."""
    import socket
    import time
    import urllib.request
    port = free_port()
    env = dict(os.environ, B200_DISABLE_KERNELS="1", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, "tools", "run_text_generation_server.py")] + model_args + extra + [
        "--host", "127.0.0.1", "--port", str(port)]
    log_path = tmp_path / f"server_{tag}.log"
    log = open(log_path, "w")
    server = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, env=env, cwd=ROOT)
    try:
        deadline = time.time() + 180
        while True:
            assert server.poll() is None and time.time() < deadline, open(log_path).read()[-3000:]
            try:
                socket.create_connection(("127.0.0.1", port), timeout=1).close()
                break
            except OSError:
                time.sleep(1)
        answers = []
        for body in bodies:
            req = urllib.request.Request(f"http://127.0.0.1:{port}/api", method="PUT", data=json.dumps(body).encode(),
                                         headers={"Content-Type": "application/json"})
            with urllib.request.urlopen(req, timeout=180) as resp:
                answers.append(json.loads(resp.read().decode()))
        return answers
    finally:
        server.terminate()
        try:
            server.wait(timeout=30)
        except subprocess.TimeoutExpired:
            server.kill()
        log.close()


def test_pipeline_parallel_server_matches_single_rank(tmp_path):
    """The same trained weights served from one rank and, resharded by tools/checkpoint_util.py, from two
    pipeline stages: greedy sampling with log-probabilities, beam search and prompt scoring return
    identical text and log-probabilities (megatron/text_generation/forward_step.py pipelined path)."""
    from tests.test_cli_e2e_cpu import _corpus
    _corpus(str(tmp_path / "corpus"))
    model = ["--model_name", "llama2", "--num_layers", "4", "--hidden_size", "64", "--num_attention_heads", "4",
             "--num_attention_heads_kv", "2", "--ffn_hidden_size", "128", "--seq_length", "32",
             "--max_position_embeddings", "32", "--micro_batch_size", "2", "--use_rms_norm", "--glu_activation", "swiglu",
             "--no_tie_embed_logits", "--position_embedding_type", "rotary", "--no_bias_gelu_fusion",
             "--hidden_dropout", "0", "--attention_dropout", "0", "--tokenizer_type", "NullTokenizer",
             "--vocab_size", "1021", "--distributed_backend", "gloo", "--use_cpu_initialization"]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", B200_DISABLE_KERNELS="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT=str(free_port()), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "finetune.py")] + model + [
        "--global_batch_size", "4", "--lr", "1e-3", "--train_iters", "2", "--data_path", str(tmp_path / "corpus"),
        "--split", "90,5,5", "--log_interval", "1", "--eval_iters", "0", "--num_workers", "0",
        "--save", str(tmp_path / "ck"), "--save_interval", "2"], capture_output=True, text=True, env=env, cwd=ROOT,
        timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "checkpoint_util.py"), "--model_type", "llama2",
                        "--load_dir", str(tmp_path / "ck"), "--save_dir", str(tmp_path / "ck_pp2"),
                        "--target_tensor_parallel_size", "1", "--target_pipeline_parallel_size", "2"],
                       capture_output=True, text=True, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    bodies = [{"prompts": ["5 6 7 8", "9 10"], "tokens_to_generate": 6, "top_k": 1, "logprobs": True},
              {"prompts": ["5 6 7 8"], "tokens_to_generate": 4, "beam_width": 2},
              {"prompts": ["5 6 7 8 9 10 11"], "tokens_to_generate": 0, "logprobs": True}]
    load = ["--no_load_optim", "--no_load_rng", "--load"]
    one = _serve_and_query(tmp_path, "pp1", 1, model, load + [str(tmp_path / "ck")], bodies)
    two = _serve_and_query(tmp_path, "pp2", 2, model,
                           ["--pipeline_model_parallel_size", "2"] + load + [str(tmp_path / "ck_pp2")], bodies)
    assert len(one[0]["text"][0].split()) == 10 and len(one[1]["text"]) == 2
    for a, b in zip(one, two):
        assert a["text"] == b["text"]
        if a.get("logprobs"):
            flat_a, flat_b = torch.tensor(a["logprobs"][0]), torch.tensor(b["logprobs"][0])
            assert torch.allclose(flat_a, flat_b, atol=1e-5)
