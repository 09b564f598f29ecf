"""Tensor-parallel primitives on CPU/Gloo: the 7 mappings fwd+bwd, vocab-parallel CE vs
F.cross_entropy, broadcast_data, RNG tracker + checkpoint, 1-D split/gather, layers vs a
dense master weight (ideas of the reference's test families A and C)."""
import pytest
import torch
import torch.nn.functional as F

from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.parallel import (mappings, cross_entropy, data as tp_data, random as tp_random,
                                         utils as tp_utils, layers)
from tests.helpers.dist import spawn


def _mappings(rank, world):
    tp = mpu.get_tensor_model_parallel_world_size()
    r = mpu.get_tensor_model_parallel_rank()
    # copy <-> reduce
    x = torch.full((4, 6), float(r + 1), requires_grad=True)
    y = mappings.copy_to_tensor_model_parallel_region(x)
    assert torch.equal(y, x)
    y.sum().backward()
    assert torch.allclose(x.grad, torch.full_like(x, float(tp)))
    x = torch.full((4, 6), float(r + 1), requires_grad=True)
    y = mappings.reduce_from_tensor_model_parallel_region(x * 1.0)
    assert torch.allclose(y, torch.full_like(y, sum(range(1, tp + 1))))
    y.sum().backward()
    assert torch.allclose(x.grad, torch.ones_like(x))
    # scatter / gather along the last dim
    full = torch.arange(4 * 2 * tp, dtype=torch.float).view(4, 2 * tp).requires_grad_()
    part = mappings.scatter_to_tensor_model_parallel_region(full)
    assert torch.equal(part, full[:, 2 * r:2 * r + 2])
    back = mappings.gather_from_tensor_model_parallel_region(part)
    assert torch.equal(back, full)
    back.sum().backward()
    assert torch.allclose(full.grad, torch.ones_like(full))
    # sequence-parallel scatter / gather / reduce-scatter along the first dim
    seq = torch.arange(2 * tp * 3, dtype=torch.float).view(2 * tp, 3).requires_grad_()
    sp = mappings.scatter_to_sequence_parallel_region(seq)
    assert torch.equal(sp, seq[2 * r:2 * r + 2])
    g = mappings.gather_from_sequence_parallel_region(sp, True)
    assert torch.equal(g, seq)
    g.sum().backward()            # backward of gather = reduce-scatter of ones -> tp
    assert torch.allclose(seq.grad[2 * r:2 * r + 2], torch.full((2, 3), float(tp)))
    z = torch.full((2 * tp, 3), float(r + 1), requires_grad=True)
    rs = mappings.reduce_scatter_to_sequence_parallel_region(z * 1.0)
    assert rs.shape == (2, 3) and torch.allclose(rs, torch.full_like(rs, sum(range(1, tp + 1))))
    rs.sum().backward()
    assert torch.allclose(z.grad, torch.ones_like(z))


@pytest.mark.parametrize("world", [1, 2, 4])
def test_mappings(world):
    spawn(_mappings, world, tp=world)


def _cross_entropy(rank, world, smoothing):
    tp, r = mpu.get_tensor_model_parallel_world_size(), mpu.get_tensor_model_parallel_rank()
    torch.manual_seed(3)
    V, T = 16 * tp, 12
    logits = torch.randn(3, 4, V)
    target = torch.randint(0, V, (3, 4))
    ref_in = logits.clone().requires_grad_()
    ref = F.cross_entropy(ref_in.view(-1, V), target.view(-1), reduction="none",
                          label_smoothing=0.0).view(3, 4)
    mine_in = logits[..., r * 16:(r + 1) * 16].clone().requires_grad_()
    loss = cross_entropy.vocab_parallel_cross_entropy(mine_in, target, smoothing)
    if smoothing == 0.0:
        assert torch.allclose(loss, ref, atol=1e-5)
        w = torch.rand(3, 4)
        (loss * w).sum().backward()
        (ref * w).sum().backward()
        assert torch.allclose(mine_in.grad, ref_in.grad[..., r * 16:(r + 1) * 16], atol=1e-5)
    else:
        assert torch.isfinite(loss).all() and (loss > 0).all()
        loss.sum().backward()
        assert torch.isfinite(mine_in.grad).all()


@pytest.mark.parametrize("world", [1, 2])
@pytest.mark.parametrize("smoothing", [0.0, 0.1])
def test_vocab_parallel_cross_entropy(world, smoothing):
    spawn(_cross_entropy, world, smoothing, tp=world)


def _broadcast_data(rank, world):
    keys = ["a", "b"]
    data = None
    if mpu.get_tensor_model_parallel_rank() == 0:
        data = {"a": torch.arange(12).view(3, 4), "b": torch.arange(5) + 100}
    out = tp_data.broadcast_data(keys, data, torch.int64)
    assert torch.equal(out["a"].cpu(), torch.arange(12).view(3, 4))
    assert torch.equal(out["b"].cpu(), torch.arange(5) + 100)


@pytest.mark.parametrize("world", [1, 2])
def test_broadcast_data(world):
    spawn(_broadcast_data, world, tp=world)


def _rng_and_checkpoint(rank, world):
    tracker = tp_random.get_cuda_rng_tracker()
    with tracker.fork():
        a = torch.rand(4)
    outs = [torch.zeros(4) for _ in range(world)]
    torch.distributed.all_gather(outs, a)
    if world > 1:
        assert not torch.equal(outs[0], outs[1])          # TP stream differs per rank
    b = torch.rand(4)                                     # default stream identical on all ranks
    outs = [torch.zeros(4) for _ in range(world)]
    torch.distributed.all_gather(outs, b)
    assert all(torch.equal(outs[0], o) for o in outs)
    with pytest.raises(Exception):
        tracker.add("model-parallel-rng", 999)
    # checkpoint: same gradients as the plain function, dropout mask replayed
    lin = torch.nn.Linear(8, 8)
    x = torch.randn(2 * world, 8, requires_grad=True)

    def fn(t):
        with tracker.fork():
            return F.dropout(lin(t), 0.5, training=True).sum()
    states = tracker.get_states()
    ref = fn(x)
    ref.backward()
    g_ref, gw_ref = x.grad.clone(), lin.weight.grad.clone()
    x.grad = None
    lin.weight.grad = None
    tracker.set_states(states)
    out = tp_random.checkpoint(fn, False, x)
    out.backward()
    assert torch.allclose(out, ref) and torch.allclose(x.grad, g_ref) and torch.allclose(lin.weight.grad, gw_ref)


@pytest.mark.parametrize("world", [1, 2])
def test_rng_tracker_and_checkpoint(world):
    spawn(_rng_and_checkpoint, world, tp=world)


def _split_gather(rank, world):
    tp, r = mpu.get_tensor_model_parallel_world_size(), mpu.get_tensor_model_parallel_rank()
    t = torch.arange(4 * tp * 3, dtype=torch.float).view(4 * tp, 3)
    part = tp_utils.split_tensor_into_1d_equal_chunks(t)
    assert torch.equal(part, t.view(-1)[r * 12:(r + 1) * 12])
    assert torch.equal(tp_utils.gather_split_1d_tensor(part), t.view(-1))
    assert tp_utils.VocabUtility.vocab_range_from_global_vocab_size(8 * tp, r, tp) == (8 * r, 8 * r + 8)
    parts = tp_utils.split_tensor_along_last_dim(torch.zeros(2, 6), 3)
    assert len(parts) == 3 and parts[0].shape == (2, 2)
    assert tp_utils.divide(12, 4) == 3
    with pytest.raises(AssertionError):
        tp_utils.divide(12, 5)


@pytest.mark.parametrize("world", [1, 2])
def test_split_gather_utils(world):
    spawn(_split_gather, world, tp=world)


def _layers_vs_dense(rank, world, sequence_parallel):
    tp, r = mpu.get_tensor_model_parallel_world_size(), mpu.get_tensor_model_parallel_rank()
    torch.manual_seed(11)
    S, B, H, O = 4 * tp, 2, 16, 8 * tp
    init = lambda w: torch.nn.init.normal_(w, std=0.1)   # noqa: E731
    torch.manual_seed(5)
    col = layers.ColumnParallelLinear(H, O, bias=True, gather_output=False, init_method=init,
                                      use_cpu_initialization=True, keep_master_weight_for_test=True,
                                      sequence_parallel_enabled=sequence_parallel,
                                      async_tensor_model_parallel_allreduce=False)
    row = layers.RowParallelLinear(O, H, bias=True, input_is_parallel=True, init_method=init,
                                   use_cpu_initialization=True, keep_master_weight_for_test=True,
                                   sequence_parallel_enabled=sequence_parallel)
    Wc, Wr = col.master_weight.float(), row.master_weight.float()
    torch.manual_seed(21)
    x_full = torch.randn(S, B, H)
    x_ref = x_full.clone().requires_grad_()
    ref = F.linear(F.gelu(F.linear(x_ref, Wc)), Wr)
    ref.pow(2).sum().backward()
    if sequence_parallel and tp > 1:
        x = x_full[r * 4:(r + 1) * 4].clone().requires_grad_()
    else:
        x = x_full.clone().requires_grad_()
    h, _ = col(x)
    y, _ = row(F.gelu(h))
    if sequence_parallel and tp > 1:
        assert torch.allclose(y, ref[r * 4:(r + 1) * 4], atol=1e-5)
        # each rank differentiates its own shard of the loss (sum over shards == full loss)
        y.pow(2).sum().backward()
        assert torch.allclose(x.grad, x_ref.grad[r * 4:(r + 1) * 4], atol=1e-4)
    else:
        assert torch.allclose(y, ref, atol=1e-5)
        y.pow(2).sum().backward()
        assert torch.allclose(x.grad, x_ref.grad, atol=1e-4)
    # weight gradients equal the matching shard of the dense gradient
    dWc = torch.autograd.grad(F.linear(F.gelu(F.linear(x_full, Wc.requires_grad_())), Wr).pow(2).sum(), Wc)[0]
    n = O // tp
    assert torch.allclose(col.weight.grad, dWc[r * n:(r + 1) * n], atol=1e-4)
    # embedding
    torch.manual_seed(9)
    emb = layers.VocabParallelEmbedding(8 * tp, H, init_method=init, use_cpu_initialization=True)
    ids = torch.randint(0, 8 * tp, (B, S))
    out = emb(ids)
    torch.manual_seed(9)
    master = torch.empty(8 * tp, H)
    init(master)
    assert torch.allclose(out, F.embedding(ids, master), atol=1e-6)


@pytest.mark.parametrize("world,sp", [(1, False), (2, False), (2, True)])
def test_parallel_layers_match_dense(world, sp):
    spawn(_layers_vs_dense, world, sp, tp=world)
