"""Pipeline and tensor parallel schedules on CPU / Gloo (world size 2): 1F1B, interleaved 1F1B with
two stages (the 4-way exchange talks to ONE peer in both directions, so the posting order of the
sends and receives decides which tensor lands where) and TP + sequence parallelism must all follow
the single-process loss curve of the same tiny Llama."""
import pytest

from tests.helpers.launch import run_tiny

KW = dict(steps=3, layers=4, num_microbatches=4, backend="gloo", dtype="fp32", hidden=128, heads=4, kv_heads=2,
          seq=64)
ENV = {"B200_DISABLE_KERNELS": "1", "CUDA_VISIBLE_DEVICES": ""}


@pytest.fixture(scope="module")
def base():
    return run_tiny(env_extra=ENV, **KW)["losses"]


def _close(a, b, rel):
    assert len(a) == len(b) > 0
    for x, y in zip(a, b):
        assert abs(x - y) <= rel * abs(y), (a, b)


def test_pp2_1f1b(base):
    _close(run_tiny(nproc=2, pp=2, env_extra=ENV, **KW)["losses"], base, 2e-3)


def test_pp2_interleaved(base):
    _close(run_tiny(nproc=2, pp=2, vpp_layers=1, env_extra=ENV, **KW)["losses"], base, 2e-3)


def test_tp2_sequence_parallel(base):
    _close(run_tiny(nproc=2, tp=2, env_extra=ENV, **KW)["losses"], base, 2e-3)


def test_dp2_overlapped_bucketed_grad_reduce_matches_single_reduce():
    """Gradients averaged bucket by bucket during backward (many small buckets) == one reduction
    after backward, for the plain and the distributed optimizer."""
    kw = dict(KW, num_microbatches=2, layers=2)
    ref = run_tiny(nproc=2, no_overlap=True, env_extra=ENV, **kw)["losses"]
    for extra in (dict(bucket_mb=0), dict(bucket_mb=0, dist_opt=True)):
        got = run_tiny(nproc=2, env_extra=ENV, **kw, **extra)["losses"]
        _close(got, ref, 1e-5)


def test_dp2_tied_embeddings_overlap_counts_every_use():
    """A tied GPT (word embedding == LM head, pp = 1) writes the embedding gradient twice per
    micro-batch: the head's weight-gradient epilogue (outside autograd) and the lookup gradient
    (autograd hook). The bucket holding it must not be reduced before both have landed in the last
    micro-batch: overlapped == one reduction after backward, with the weight-gradient epilogue
    restated on the host so the arrival order is the GPU's."""
    kw = dict(KW, model_name="gpt", num_microbatches=2, layers=2, steps=4)
    env = dict(ENV, B200_HOST_FUSED_WGRAD="1")
    ref = run_tiny(nproc=2, no_overlap=True, env_extra=env, **kw)
    for extra in (dict(bucket_mb=0), dict(bucket_mb=0, dist_opt=True), dict(bucket_mb=0, num_microbatches=1)):
        k2 = dict(kw, **extra)
        want = ref if "num_microbatches" not in extra else run_tiny(nproc=2, no_overlap=True, env_extra=env, **k2)
        got = run_tiny(nproc=2, env_extra=env, **k2)
        _close(got["losses"], want["losses"], 1e-5)
        _close(got["grad_norms"], want["grad_norms"], 1e-5)


def test_dp2_ddp_impl_torch_matches_local():
    """--DDP_impl torch (parameter-dtype gradients averaged in 25 MiB buckets during backward) trains
    like --DDP_impl local with one reduction after backward."""
    kw = dict(KW, num_microbatches=2, layers=2)
    ref = run_tiny(nproc=2, no_overlap=True, env_extra=ENV, **kw)["losses"]
    got = run_tiny(nproc=2, extra="--DDP_impl=torch", env_extra=ENV, **kw)["losses"]
    _close(got, ref, 1e-5)


def test_dp_consistency_check_passes_and_catches_divergence():
    """--check_dp_consistency_interval: parameter checksums agree across replicas after every step;
    a replica whose weights were perturbed is reported."""
    kw = dict(KW, num_microbatches=2, layers=2, dist_opt=True, check_dp=1)
    assert len(run_tiny(nproc=2, env_extra=ENV, **kw)["losses"]) == 3
    # (under the distributed optimizer a perturbed replica is healed by the next parameter
    # all-gather; the replicated optimizer keeps it)
    with pytest.raises(AssertionError, match="replicas diverged"):
        run_tiny(nproc=2, env_extra=ENV, corrupt_rank=1, **dict(kw, dist_opt=False))


def test_falcon_gqa_tp2_matches_single_process():
    """Falcon block (parallel attention + MLP, two LayerNorms, GeLU, tied embeddings) with grouped KV
    heads sharded over two tensor-parallel ranks under sequence parallelism."""
    kw = dict(KW, model_name="falcon", layers=2, num_microbatches=2)
    one = run_tiny(env_extra=ENV, **kw)["losses"]
    _close(run_tiny(nproc=2, tp=2, env_extra=ENV, **kw)["losses"], one, 2e-3)


def test_tp2_without_sequence_parallel_and_distributed_saved_activations(base):
    """Plain tensor parallelism (all-reduce instead of all-gather / reduce-scatter) under full
    recompute with the checkpointed layer inputs sharded over the TP ranks
    (--distribute_saved_activations, megatron/core/tensor_parallel/random.py:198-218): the CPU
    initialisation is TP-invariant, so the losses match the single-process run to rounding."""
    got = run_tiny(nproc=2, tp=2, no_sp=True, recompute=True, extra="--distribute_saved_activations",
                   env_extra=ENV, **KW)["losses"]
    _close(got, base, 1e-5)


def test_pp2_standalone_embedding_stage(base):
    """--standalone_embedding_stage: the first pipeline rank holds only the embedding, the
    transformer layers all live on the second."""
    got = run_tiny(nproc=2, pp=2, extra="--standalone_embedding_stage", env_extra=ENV, **KW)["losses"]
    _close(got, base, 2e-3)
    assert got[-1] < got[0]
