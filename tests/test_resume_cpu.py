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
