"""Host-side bookkeeping of the tensor-parallel peer-memory workspace (parallel/symm.py) without a GPU.

The kernels' cross-GPU protocol rests on numbers the Python side hands them: the epoch sequence shared by
the all-gather GEMM and the stand-alone staging windows, which staging buffer / counter set a call uses,
which counter set it must leave zeroed for the next call, and the running targets of the completion
counters. A call sequence that mixes both users of the staging buffers once left the all-gather GEMM
waiting for a counter value that could never be reached; this test replays such sequences against a
recording stand-in for the native library and checks every number."""
import ctypes as C

import pytest
import torch

from epfl_megatron_b200.ops import lib as L
from epfl_megatron_b200.parallel import symm


class _Recorder:
    def __init__(self):
        self.calls = []
        self.next_ptr = 0x10000000

    def call(self, name, *args, launches=1):
        if name == "b200_symm_alloc":
            args[1]._obj.value = self.next_ptr           # C.byref(c_void_p)
            self.next_ptr += 0x10000000
        elif name == "b200_symm_import":
            args[1]._obj.value = self.next_ptr
            self.next_ptr += 0x10000000
        self.calls.append((name, args))
        return 0

    def last(self, name):
        return [a for n, a in self.calls if n == name][-1]


@pytest.fixture
def ws(monkeypatch):
    try:
        lib = L.lib()
    except Exception:  # noqa: BLE001
        pytest.skip("native library not built")
    rec = _Recorder()
    monkeypatch.setattr(symm.L, "call", rec.call)
    monkeypatch.setattr(symm.L, "stream_ptr", lambda: 0)
    monkeypatch.setattr(symm.dist, "get_world_size", lambda group=None: 2)
    monkeypatch.setattr(symm.dist, "get_rank", lambda group=None: 1)
    monkeypatch.setattr(symm.dist, "barrier", lambda group=None: None)

    def gather(out, obj, group=None):
        for i in range(len(out)):
            out[i] = obj if i == 1 else bytes(64)
    monkeypatch.setattr(symm.dist, "all_gather_object", gather)
    monkeypatch.setattr(symm.torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(symm, "_wrap", lambda ptr, nbytes, device: torch.zeros(nbytes, dtype=torch.uint8))
    w = symm.TPWorkspace(group=None, max_rows=1024, max_k=512, max_n=512)
    w.device = torch.device("cpu")
    w.rec, w.lib = rec, lib
    return w


def _desc_stage(ws, desc_ref):
    """Which staging buffer (0 / 1) the descriptor passed to the GEMM points at."""
    d = desc_ref._obj
    return 0 if d.peer_stage[ws.rank] == ws._stage(ws.rank, 0) else 1


def test_epochs_counter_sets_and_targets_across_mixed_calls(ws):
    mb = ws.max_blocks
    x = torch.zeros(512, 512, dtype=torch.bfloat16)
    wgt = torch.zeros(256, 512, dtype=torch.bfloat16)
    seen = []
    for step, kind in enumerate(["gemm", "gemm", "window", "gemm", "window", "window", "gemm"], start=1):
        ws._flags.fill_(7)                                  # whatever the previous calls left behind
        if kind == "gemm":
            ws.allgather_gemm(x, wgt)
            a = ws.rec.last("b200_gemm_allgather")
            epoch, stage_call = a[11], a[13]
            assert epoch == step, "one epoch per user of the staging buffers, whoever it is"
            assert _desc_stage(ws, a[0]) == step % 2, "staging buffers alternate with the epoch"
            assert a[12] == ws._flag(ws.rank, ws.w_stage_done)
            seen.append(stage_call)
            # the kernel itself clears the other counter set; the host must not have touched the flags
            assert int(ws._flags.min()) == 7
        else:
            e, srcs = ws.gather_begin(x)
            assert e == step
            st = ws.rec.last("b200_stage_shard")
            assert st[1] == ws._stage(ws.rank, step % 2) and st[3] == 512 * 512 * 2
            assert st[7] == (step - 2) & 0xFFFFFFFF and st[8] == int(step > 2), "waits for the pull of epoch e - 2"
            assert st[9] == ws._flag(ws.rank, ws.w_stage_done2), "own completion counter, not the GEMM's"
            assert st[10] == ws.stage_calls and st[13] == step
            wt = ws.rec.last("b200_symm_wait")
            assert wt[0] == ws._flag(ws.rank, ws.w_ready) and wt[1:4] == (ws.tp, ws.rank, step)
            assert srcs[ws.rank] == x.data_ptr() and srcs[0] == ws._stage(0, step % 2)
            # the all-gather GEMM of epoch e + 1 finds ITS counter set zeroed, nothing else was touched
            nxt = (step + 1) & 1
            flags = ws._flags
            assert int(flags[nxt * mb:(nxt + 1) * mb].abs().sum()) == 0
            assert int(flags[2 * mb + 8 + nxt * 8:2 * mb + 16 + nxt * 8].abs().sum()) == 0
            keep = torch.ones_like(flags, dtype=torch.bool)
            keep[nxt * mb:(nxt + 1) * mb] = False
            keep[2 * mb + 8 + nxt * 8:2 * mb + 16 + nxt * 8] = False
            assert bool((flags[keep] == 7).all())
            ws.gather_end(e)
            sg = ws.rec.last("b200_symm_signal")
            assert sg[1] == 1 and sg[2] == step
            assert sg[0][0] == ws._flag(0, ws.w_consumed + ws.rank), "peer 0's consumed[<me>] word"
    # the GEMM's staging counter advances once per GEMM call, not once per epoch
    assert seen == [1, 2, 3, 4]
    assert ws.ag_epoch == 7 and ws.stage_calls == 3 and ws.ag_gemm_calls == 4


def test_reduce_sets_alternate_and_cross_entropy_targets_accumulate(ws):
    a = torch.zeros(1024, 256, dtype=torch.bfloat16)
    wgt = torch.zeros(512, 256, dtype=torch.bfloat16)
    sets = []
    for _ in range(4):
        ws.gemm_reduce_scatter(a, wgt)
        r = ws.rec.last("b200_gemm_reduce_scatter")
        sets.append(r[2])
        assert r[1][ws.rank] == ws._parts(ws.rank, r[2]) and r[1][0] == ws._parts(0, r[2])
    assert sets == [1, 0, 1, 0]
    logits = torch.zeros(300, 64, dtype=torch.bfloat16)
    labels = torch.zeros(300, dtype=torch.int64)
    grid = ws.lib.b200_xent_push_grid(300)
    targets = []
    for _ in range(4):
        ws.xent_exchange(logits, labels, 0)
        push, comb = ws.rec.last("b200_xent_stats_push"), ws.rec.last("b200_xent_combine_wait")
        s = ws.xent_calls % 2
        assert push[2][0] == ws._xent(0, s) and push[3][0] == ws._flag(0, ws.w_xent_cnt + s)
        assert comb[0] == ws._xent(ws.rank, s) and comb[1] == ws._flag(ws.rank, ws.w_xent_cnt + s)
        targets.append((s, comb[2]))
    # every call adds tp * grid arrivals to the counter of ITS set
    assert targets == [(1, 2 * grid), (0, 2 * grid), (1, 4 * grid), (0, 4 * grid)]


def test_flag_words_do_not_overlap(ws):
    words = ws.lib.b200_comm_flag_words(ws.max_blocks, ws.max_tn)
    extra = [ws.w_stage_done, ws.w_xent_cnt, ws.w_xent_cnt + 1, ws.w_stage_done2] + \
            [ws.w_table_ready + i for i in range(8)]
    assert min(extra) >= words and len(set(extra)) == len(extra)
    assert (max(extra) + 1) * 4 <= ws.flag_bytes
    assert ws.w_ready == 2 * ws.max_blocks and ws.w_consumed == 2 * ws.max_blocks + 24
