"""Pipeline schedule simulator: every pipeline rank's schedule is run
in this one process against a recording stand-in for torch.distributed's batched send / receive.
The recorded per-rank sequences of posted sends and receives are then replayed against each other
with NCCL's matching rule (between one pair of ranks, sends and receives match in posting order; a
batch completes when all of its operations are matched):

* the schedules must run to completion on every rank (no deadlock, nothing left unmatched),
* every receive of an activation must be matched by a send of an activation (and gradients by
  gradients) -- the two-stage interleaved case, where the neighbour in both directions is the same
  rank, is exactly where the posting order decides that,
* 1F1B: rank r runs pp - r - 1 warm-up forwards, and never holds more than pp - r micro-batches.

Reference: megatron/schedules.py:253-722, megatron/p2p_communication.py:101-251.
"""
import types

import pytest
import torch

from epfl_megatron_b200 import global_vars, p2p_communication, schedules
from epfl_megatron_b200.models.enums import ModelType
from epfl_megatron_b200.parallel import state as mpu

H, S, MB = 8, 4, 1
_REAL_COMMUNICATE = p2p_communication._communicate
_REAL_BACKWARD_STEP = schedules.backward_step


@pytest.fixture
def sim_state():
    """The simulator drives the rank / size overrides of parallel.state: put them back afterwards."""
    yield
    for setter in (mpu.set_pipeline_model_parallel_world_size, mpu.set_pipeline_model_parallel_rank,
                   mpu.set_tensor_model_parallel_world_size, mpu.set_tensor_model_parallel_rank,
                   mpu.set_virtual_pipeline_model_parallel_world_size, mpu.set_virtual_pipeline_model_parallel_rank):
        setter(None)


class _Req:
    def wait(self):
        return None


class _FakeDist:
    """Stands in for ``torch.distributed`` inside p2p_communication: records what is posted."""

    def __init__(self, log):
        self.log = log

    @staticmethod
    def isend(*a, **k):
        raise AssertionError("unbatched send")

    @staticmethod
    def irecv(*a, **k):
        raise AssertionError("unbatched recv")

    def P2POp(self, op, tensor, peer, group=None):
        return types.SimpleNamespace(kind="send" if op is self.isend else "recv", tensor=tensor, peer=peer)

    def batch_isend_irecv(self, ops):
        batch = []
        for o in ops:
            if o.kind == "recv":
                o.tensor.data.zero_()
            # tag: what the SENDER means / what the RECEIVER expects, from the direction
            batch.append((o.kind, o.peer, getattr(o.tensor, "_sim_tag", None) if o.kind == "send" else id(o.tensor)))
        self.log.append(batch)
        return [_Req() for _ in ops]

    @staticmethod
    def get_backend():
        return "nccl"


class _Stage(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.w = torch.nn.Parameter(torch.ones(H))
        self.inp = None

    def set_input_tensor(self, t):
        self.inp = t[0] if isinstance(t, list) else t

    def forward(self):
        x = self.inp if self.inp is not None else torch.ones(S, MB, H)
        return x * self.w + 1.0


def _run_rank(monkeypatch, pp, rank, n_mb, n_chunks, events):
    """Run the schedule of one pipeline rank; returns its list of posted batches."""
    log = []
    fake = _FakeDist(log)
    monkeypatch.setattr(p2p_communication, "dist", fake)
    monkeypatch.setattr(p2p_communication, "_device", lambda: torch.device("cpu"))
    args = types.SimpleNamespace(seq_length=S, micro_batch_size=MB, hidden_size=H, sequence_parallel=False,
                                 variable_seq_lengths=False, scatter_gather_tensors_in_pipeline=False,
                                 params_dtype=torch.float32, fp32_residual_connection=False,
                                 model_type=ModelType.encoder_or_decoder, decoder_seq_length=None,
                                 pipeline_model_parallel_size=pp,
                                 virtual_pipeline_model_parallel_size=n_chunks if n_chunks > 1 else None)
    monkeypatch.setattr(global_vars, "get_args", lambda: args)
    monkeypatch.setattr(global_vars, "get_num_microbatches", lambda: n_mb)
    monkeypatch.setattr(mpu, "_PP_RANKS", list(range(pp)), raising=False)
    mpu.set_pipeline_model_parallel_world_size(pp)
    mpu.set_pipeline_model_parallel_rank(rank)
    mpu.set_tensor_model_parallel_world_size(1)
    mpu.set_tensor_model_parallel_rank(0)
    mpu.set_virtual_pipeline_model_parallel_world_size(n_chunks if n_chunks > 1 else None)
    mpu.set_virtual_pipeline_model_parallel_rank(0 if n_chunks > 1 else None)

    # tag what travels: _communicate sends output tensors "next" and gradients "prev"
    real = _REAL_COMMUNICATE

    def tagged(send_next, send_prev, recv_prev, recv_next, *a, **k):
        if send_next is not None:
            send_next = send_next.contiguous()
            send_next._sim_tag = "act"
        if send_prev is not None:
            send_prev = send_prev.contiguous()
            send_prev._sim_tag = "grad"
        before = len(log)
        out = real(send_next, send_prev, recv_prev, recv_next, *a, **k)
        if len(log) > before:
            # receives were logged by buffer identity: the buffer returned as "from prev" carries an
            # activation, the one returned as "from next" a gradient -- in whatever order they were posted
            kinds = {id(out[0]): "act", id(out[1]): "grad"}
            log[-1] = [(kind, peer, kinds[tag] if kind == "recv" else tag) for kind, peer, tag in log[-1]]
        return out

    monkeypatch.setattr(p2p_communication, "_communicate", tagged)

    def fwd_func(data_iterator, model):
        events.append(("F", rank))
        out = model()

        def loss_func(o):
            loss = o.sum()
            return loss, {"loss": loss.detach()}
        return out, loss_func

    real_bwd = _REAL_BACKWARD_STEP

    def counted_bwd(*a, **k):
        events.append(("B", rank))
        return real_bwd(*a, **k)

    monkeypatch.setattr(schedules, "backward_step", counted_bwd)
    models = [_Stage() for _ in range(n_chunks)]
    fb = (schedules.forward_backward_pipelining_with_interleaving if n_chunks > 1
          else schedules.forward_backward_pipelining_without_interleaving)
    data = [None] * n_chunks if n_chunks > 1 else None
    fb(fwd_func, data, models, None, None, forward_only=False)
    for m in models:
        assert m.w.grad is not None
    return log


def _replay(logs):
    """NCCL matching between the ranks' posted batches; returns the number of matched messages."""
    pp = len(logs)
    pos = [0] * pp                                   # current batch of every rank
    pending = [None] * pp                            # unmatched operations of the current batch

    def load(r):
        pending[r] = list(logs[r][pos[r]]) if pos[r] < len(logs[r]) else None

    for r in range(pp):
        load(r)
    matched = 0
    while any(p is not None for p in pending):
        progress = False
        for a in range(pp):
            if pending[a] is None:
                continue
            for op in list(pending[a]):
                kind, b, tag = op
                if kind != "send" or pending[b] is None:
                    continue
                # first still-unmatched send a->b of a's batch, first unmatched recv from a of b's batch
                if op is not next(o for o in pending[a] if o[0] == "send" and o[1] == b):
                    continue
                recv = next((o for o in pending[b] if o[0] == "recv" and o[1] == a), None)
                if recv is None:
                    continue
                assert recv[2] == tag, f"rank {b} expected a {recv[2]} from rank {a} but the matching send is a {tag}"
                pending[a].remove(op)
                pending[b].remove(recv)
                matched += 1
                progress = True
            for r in (a,) + tuple(range(pp)):
                if pending[r] is not None and not pending[r]:
                    pos[r] += 1
                    load(r)
                    progress = True
        assert progress, f"deadlock: batches {pos} of {[len(l) for l in logs]}, pending {pending}"
    assert all(pos[r] == len(logs[r]) for r in range(pp))
    return matched


def _simulate(monkeypatch, pp, n_mb, n_chunks):
    events, logs = [], []
    for rank in range(pp):
        logs.append(_run_rank(monkeypatch, pp, rank, n_mb, n_chunks, events))
    return logs, events


@pytest.mark.parametrize("pp,n_mb", [(2, 4), (4, 8), (4, 2), (3, 5)])
def test_1f1b_schedule(monkeypatch, sim_state, pp, n_mb):
    logs, events = _simulate(monkeypatch, pp, n_mb, 1)
    # one activation forward and one gradient back per micro-batch and stage boundary
    assert _replay(logs) == 2 * n_mb * (pp - 1)
    for r in range(pp):
        seq = [k for k, rk in events if rk == r]
        assert seq.count("F") == seq.count("B") == n_mb
        warm = min(pp - r - 1, n_mb)
        lead = min(warm + 1, n_mb)                   # warm-up forwards + the first steady-state forward
        assert seq[:lead] == ["F"] * lead and seq[lead] == "B"
        live = peak = 0
        for k in seq:
            live += 1 if k == "F" else -1
            peak = max(peak, live)
        assert peak == min(pp - r, n_mb)


@pytest.mark.parametrize("pp,n_mb,chunks", [(2, 4, 2), (2, 2, 2), (4, 8, 2), (4, 4, 3), (2, 6, 3)])
def test_interleaved_schedule(monkeypatch, sim_state, pp, n_mb, chunks):
    logs, events = _simulate(monkeypatch, pp, n_mb, chunks)
    # every (micro-batch, virtual stage) boundary carries one activation and one gradient; the ring
    # closes from the last rank to the first between consecutive chunks
    assert _replay(logs) == 2 * n_mb * (pp * chunks - 1)
    for r in range(pp):
        seq = [k for k, rk in events if rk == r]
        assert seq.count("F") == seq.count("B") == n_mb * chunks
