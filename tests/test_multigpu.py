"""Multi-GPU parity (needs >= 2 devices; skipped otherwise): tensor + sequence parallel,
data parallel with the distributed optimizer and pipeline parallel runs must reproduce the
single-GPU loss curve of the same model (same host-side initialisation)."""
import pytest
import torch

from tests.helpers.launch import run_tiny

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _close(a, b, rel):
    assert len(a) == len(b) and len(a) > 0
    for x, y in zip(a, b):
        assert abs(x - y) <= rel * max(abs(y), 1e-3), (a, b)


@pytest.fixture(scope="module")
def single():
    return run_tiny(steps=5, num_microbatches=4)


def _fused_comm_worker(nproc):
    import json
    import os
    import subprocess
    import sys
    from tests.helpers.launch import ROOT, free_port
    port = str(free_port())
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                        "--master-addr", "127.0.0.1", "--master-port", port,
                        os.path.join(ROOT, "tests", "helpers", "fused_comm_worker.py")],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
    assert r.returncode == 0 and lines, r.stdout[-1500:] + r.stderr[-3000:]
    return json.loads(lines[-1][7:])


@pytest.mark.parametrize("nproc", [2, 4, 8])
def test_fused_collective_gemms_match_nccl(nproc):
    """All-gather -> GEMM, GEMM -> reduce-scatter and GEMM -> all-reduce in ONE kernel each over peer
    memory against NCCL + fp32 matmul: both operand layouts, repeated calls without barriers, a ragged
    output width; the all-reduce result must be bit-identical on every rank."""
    _need(nproc)
    res = _fused_comm_worker(nproc)
    for kind in ("ag", "rs", "ar"):
        assert res[kind], kind
        for rec in res[kind]:
            assert rec["max_rel_err"] < 2e-2, (kind, rec)


def test_tp2_sequence_parallel(single):
    _need(2)
    r = run_tiny(nproc=2, tp=2, steps=5, num_microbatches=4)
    _close(r["losses"], single["losses"], 0.03)


def test_tp2_without_sequence_parallel_fused_allreduce(single):
    """Plain tensor parallelism: RowParallel forward and ColumnParallel dgrad run GEMM + all-reduce in one
    kernel (parallel/layers.py _FusedRowAR / _FusedColumnAR); same loss curve as one GPU and as NCCL."""
    _need(2)
    fused = run_tiny(nproc=2, tp=2, no_sp=True, steps=5, num_microbatches=4, fused_comm="auto")
    nccl = run_tiny(nproc=2, tp=2, no_sp=True, steps=5, num_microbatches=4, fused_comm="off")
    _close(fused["losses"], single["losses"], 0.03)
    _close(fused["losses"], nccl["losses"], 0.01)


def test_tp2_tied_gpt_peer_memory_embedding_and_cross_entropy():
    """Tied GPT under TP2 + SP: the word embedding is read from the owning rank's table shard and its
    gradient accumulated into main_grad by own kernels (twice per micro-batch: lookup + LM head), the
    cross-entropy statistics travel by peer stores. Same losses as one GPU and as the NCCL path."""
    _need(2)
    kw = dict(steps=5, model_name="gpt", num_microbatches=2)
    one = run_tiny(**kw)
    fused = run_tiny(nproc=2, tp=2, **kw)
    nccl = run_tiny(nproc=2, tp=2, fused_comm="off", **kw)
    _close(fused["losses"], one["losses"], 0.03)
    _close(fused["losses"], nccl["losses"], 0.01)
    _close(fused["grad_norms"], nccl["grad_norms"], 0.05)


def test_dp2_distributed_optimizer(single):
    _need(2)
    r = run_tiny(nproc=2, steps=5, dist_opt=True, num_microbatches=2)
    assert r["losses"][-1] < r["losses"][0]


def test_dp2_fused_reduce_scatter_and_param_push():
    """In-kernel reduce-scatter + Adam-with-all-gather (parallel/dp_symm.py) against the NCCL path:
    same losses and gradient norms, with several buckets overlapped with backward."""
    _need(2)
    kw = dict(nproc=2, steps=6, dist_opt=True, num_microbatches=2, bucket_mb=1, layers=4)
    fused = run_tiny(fused_dp="on", **kw)
    nccl = run_tiny(fused_dp="off", **kw)
    assert fused["fused_dp"] and not nccl["fused_dp"]
    _close(fused["losses"], nccl["losses"], 2e-3)
    _close(fused["grad_norms"], nccl["grad_norms"], 5e-3)


def test_dp2_tied_gpt_overlap_matches_no_overlap():
    """Tied embeddings at pp = 1: the embedding gradient arrives twice per micro-batch (LM-head
    weight-gradient epilogue + lookup). Overlapped bucket reduction (NCCL and the fused path) must
    equal one reduction after backward; replicas stay identical (checksum check every step)."""
    _need(2)
    kw = dict(nproc=2, steps=5, model_name="gpt", dist_opt=True, num_microbatches=2, bucket_mb=1, layers=2,
              check_dp=1)
    ref = run_tiny(no_overlap=True, fused_dp="off", **kw)
    for mode in ("off", "on"):
        got = run_tiny(fused_dp=mode, **kw)
        _close(got["losses"], ref["losses"], 2e-3)
        _close(got["grad_norms"], ref["grad_norms"], 5e-3)


def test_dp4_fused_with_tp2():
    _need(8)
    kw = dict(nproc=8, tp=2, steps=5, dist_opt=True, num_microbatches=2, bucket_mb=1, layers=4)
    fused = run_tiny(fused_dp="on", **kw)
    nccl = run_tiny(fused_dp="off", **kw)
    assert fused["fused_dp"]
    _close(fused["losses"], nccl["losses"], 5e-3)


def test_pp2_dp2_fused_dp_matches_nccl():
    """Pipeline stages each run their own fused data-parallel reduction (one workspace per stage)."""
    _need(4)
    kw = dict(nproc=4, pp=2, steps=5, dist_opt=True, num_microbatches=4, bucket_mb=1, layers=4)
    fused = run_tiny(fused_dp="on", check_dp=1, **kw)
    nccl = run_tiny(fused_dp="off", **kw)
    assert fused["fused_dp"]
    _close(fused["losses"], nccl["losses"], 2e-3)


def test_pp2_1f1b(single):
    _need(2)
    r = run_tiny(nproc=2, pp=2, steps=5, num_microbatches=4)
    _close(r["losses"], single["losses"], 0.03)


def test_pp2_interleaved(single):
    _need(2)
    r = run_tiny(nproc=2, pp=2, vpp_layers=1, layers=4, steps=4, num_microbatches=4)
    base = run_tiny(steps=4, layers=4, num_microbatches=4)
    _close(r["losses"], base["losses"], 0.03)


def test_tp2_pp2_dp2():
    _need(8)
    r = run_tiny(nproc=8, tp=2, pp=2, steps=4, dist_opt=True, num_microbatches=4)
    assert r["losses"][-1] < r["losses"][0]
