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
