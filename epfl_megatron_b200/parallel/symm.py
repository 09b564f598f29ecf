"""Peer-mapped workspace of one tensor-parallel group and the two fused collective GEMMs on it.

Every rank of the group owns one region (csrc/comm.cu: cudaMalloc + CUDA IPC export), laid out as

    flags | 2 staging shards [rows_per_chunk, K] | 2 sets of partial-sum slots [tp][rows_per_chunk][N]
          | (only when all-reduce GEMMs are used: 2 output buffers [rows, N])

and maps the regions of its peers. ``allgather_gemm``, ``gemm_reduce_scatter`` and ``gemm_allreduce``
launch ONE kernel each (csrc/gemm_comm_tcgen05.cu) that moves activation rows over NVLink from inside the
GEMM: they replace the all-gather + matmul of ColumnParallelLinear and the matmul +
reduce-scatter of RowParallelLinear under sequence parallelism
(megatron/core/tensor_parallel/layers.py:225-243, :691-692; mappings.py:89-124).

Ordering rules the callers rely on (all ranks make the same calls in the same order):
* staging shard i is rewritten at call e only after every peer reported having pulled call e-2;
* reduce-scatter calls alternate between two slot/counter sets.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from ..ops import lib as L


class _Desc(C.Structure):
    _fields_ = [("tp", C.c_int), ("rank", C.c_int), ("peer_stage", C.c_void_p * 8),
                ("peer_flags", C.c_void_p * 8), ("max_blocks", C.c_int), ("max_tn", C.c_int)]


def _wrap(ptr: int, nbytes: int, device) -> torch.Tensor:
    """uint8 tensor over raw device memory owned by the workspace."""
    class _Mem:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                    "strides": None}
    return torch.as_tensor(_Mem(), device=device)


_OPENED = {}     # CUDA IPC handle -> mapped base address (a process opens an allocation once)


class TPWorkspace:
    """max_rows: largest M (= s*b) of a gathered activation; max_k / max_n: widest gathered input /
    reduce-scattered output in elements."""

    def __init__(self, group, max_rows: int, max_k: int, max_n: int, allreduce: bool = False):
        self.group = group
        self.tp = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert 2 <= self.tp <= 8
        self.device = torch.device("cuda", torch.cuda.current_device())
        lib = L.lib()
        self.max_rows, self.max_k, self.max_n = max_rows, max_k, max_n
        self.max_blocks = (max_rows + 127) // 128
        rpc = max_rows // self.tp
        self.max_tn = (max_n + 255) // 256
        self.allreduce = allreduce
        words = lib.b200_comm_flag_words(self.max_blocks, self.max_tn)
        self.w_ready = lib.b200_comm_shard_ready_word(self.max_blocks)
        self.w_consumed = lib.b200_comm_consumed_word(self.max_blocks)
        self.w_stage_done = words            # one extra word: completion counter of the staging warps
        self.w_xent_cnt = words + 1          # two words: arrival counters of the cross-entropy exchange
        self.w_stage_done2 = words + 3       # completion counter of the stand-alone staging kernel
        self.w_table_ready = words + 8       # [8] "rank r's embedding table is final for call n"
        self.flag_bytes = ((words + 16) * 4 + 4095) // 4096 * 4096
        self.stage_bytes = (rpc * max_k * 2 + 4095) // 4096 * 4096
        self.parts_bytes = (self.tp * rpc * max_n * 2 + 4095) // 4096 * 4096
        self.out_bytes = (max_rows * max_n * 2 + 4095) // 4096 * 4096 if allreduce else 0
        self.xent_bytes = (self.tp * max_rows * 3 * 4 + 4095) // 4096 * 4096
        self.nbytes = (self.flag_bytes + 2 * self.stage_bytes + 2 * self.parts_bytes + 2 * self.out_bytes
                       + 2 * self.xent_bytes)
        base = C.c_void_p()
        L.call("b200_symm_alloc", self.nbytes, C.byref(base), launches=0)
        self.base = base.value
        handle = (C.c_ubyte * 64)()
        L.call("b200_symm_export", self.base, handle, launches=0)
        mine = bytes(handle)
        handles = [None] * self.tp
        dist.all_gather_object(handles, mine, group=group)
        self.peer_base = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.peer_base.append(self.base)
                continue
            p = C.c_void_p()
            buf = (C.c_ubyte * 64).from_buffer_copy(h)
            L.call("b200_symm_import", buf, C.byref(p), launches=0)
            self.peer_base.append(p.value)
        dist.barrier(group=group)
        self.ag_epoch = 0
        self.rs_calls = 0
        self.xent_calls = 0
        self._xent_target = [0, 0]
        self.stage_calls = 0
        self.ag_gemm_calls = 0
        self.table_calls = 0
        self._tables = {}
        self._descs = [self._make_desc(i) for i in range(2)]
        self._flags = _wrap(self.base, self.flag_bytes, self.device).view(torch.int32)
        # Debug modes. B200_COMM_POISON=1: after every reduce-scatter the
        # partial-sum slots it consumed are overwritten with bf16 NaNs, so a later call that reads a
        # slot before its producer wrote it shows up as NaNs instead of plausible stale numbers.
        # B200_COMM_VERIFY=1: every fused call is checked against NCCL + a plain matmul (slow).
        self.poison = os.environ.get("B200_COMM_POISON", "0") == "1"
        self.verify = os.environ.get("B200_COMM_VERIFY", "0") == "1"
        self._parts_view = [
            _wrap(self._parts(self.rank, s_), self.parts_bytes, self.device).view(torch.int16) for s_ in range(2)]

    # ---- addresses -------------------------------------------------------------------------
    def _stage(self, r, i):
        return self.peer_base[r] + self.flag_bytes + i * self.stage_bytes

    def _parts(self, r, s):
        return self.peer_base[r] + self.flag_bytes + 2 * self.stage_bytes + s * self.parts_bytes

    def _out(self, r, s):
        return self.peer_base[r] + self.flag_bytes + 2 * self.stage_bytes + 2 * self.parts_bytes + s * self.out_bytes

    def _xent(self, r, s):
        return (self.peer_base[r] + self.flag_bytes + 2 * self.stage_bytes + 2 * self.parts_bytes
                + 2 * self.out_bytes + s * self.xent_bytes)

    def _flag(self, r, word):
        return self.peer_base[r] + 4 * word

    def _make_desc(self, stage):
        d = _Desc()
        d.tp, d.rank, d.max_blocks, d.max_tn = self.tp, self.rank, self.max_blocks, self.max_tn
        for r in range(self.tp):
            d.peer_stage[r] = self._stage(r, stage)
            d.peer_flags[r] = self.peer_base[r]
        return d

    # ---- all-gather -> GEMM ----------------------------------------------------------------
    def allgather_gemm(self, x_shard: torch.Tensor, weight: torch.Tensor, b_mn_major: bool = False):
        """x_shard [M/tp, K] (this rank's rows), weight [N, K] (or [K, N] when b_mn_major).
        Returns (out [M, N], gathered [M, K]); ``gathered`` is a fresh tensor the caller may keep."""
        rpc, K = x_shard.shape
        M = rpc * self.tp
        N = weight.shape[1] if b_mn_major else weight.shape[0]
        assert x_shard.is_contiguous() and x_shard.dtype == torch.bfloat16
        assert M <= self.max_rows and K <= self.max_k, (M, K, self.max_rows, self.max_k)
        self.ag_epoch += 1
        self.ag_gemm_calls += 1
        e = self.ag_epoch
        stage = e % 2
        st = L.stream_ptr()
        gathered = torch.empty(M, K, device=self.device, dtype=torch.bfloat16)
        out = torch.empty(M, N, device=self.device, dtype=torch.bfloat16)
        # ONE launch: the kernel's copier warps stage the own rows for the peers (after the peers
        # reported having pulled this staging buffer's previous content, epoch e-2), raise the
        # peers' shard_ready flags and pull the other chunks while the own rows are multiplied
        L.call("b200_gemm_allgather", C.byref(self._descs[stage]), x_shard.data_ptr(), gathered.data_ptr(),
               weight.data_ptr(), out.data_ptr(), M, N, K, int(b_mn_major), weight.stride(0), N, e,
               self._flag(self.rank, self.w_stage_done), self.ag_gemm_calls, st)
        if self.verify:
            ref = torch.empty_like(gathered)
            dist.all_gather_into_tensor(ref, x_shard, group=self.group)
            if not torch.equal(ref, gathered):
                raise RuntimeError(f"fused all-gather (call {e}, rank {self.rank}): gathered rows differ from NCCL")
            want = ref.float() @ (weight.float() if b_mn_major else weight.float().t())
            err = (out.float() - want).abs().max().item()
            if not err <= 2e-2 * max(want.abs().max().item(), 1.0):
                raise RuntimeError(f"fused all-gather GEMM (call {e}, rank {self.rank}): max abs error {err}")
        return out, gathered

    # ---- GEMM -> reduce-scatter ------------------------------------------------------------
    def gemm_reduce_scatter(self, a: torch.Tensor, weight: torch.Tensor, b_mn_major: bool = False):
        """a [M, K_local], weight [N, K_local] (or [K_local, N]); returns this rank's rows
        [M/tp, N] of the sum over ranks."""
        M, K = a.shape
        N = weight.shape[1] if b_mn_major else weight.shape[0]
        rpc = M // self.tp
        assert a.dtype == torch.bfloat16 and a.stride(1) == 1
        assert M <= self.max_rows and N <= self.max_n, (M, N, self.max_rows, self.max_n)
        self.rs_calls += 1
        s = self.rs_calls % 2
        parts = (C.c_void_p * 8)()
        for r in range(self.tp):
            parts[r] = self._parts(r, s)
        out = torch.empty(rpc, N, device=self.device, dtype=torch.bfloat16)
        L.call("b200_gemm_reduce_scatter", C.byref(self._descs[0]), parts, s, a.data_ptr(), weight.data_ptr(),
               out.data_ptr(), M, N, K, int(b_mn_major), a.stride(0), weight.stride(0), L.stream_ptr())
        if self.poison:
            self._parts_view[s].fill_(-1)       # 0xFFFF = bf16 NaN; consumed slots of this call only
        if self.verify:
            full = a.float() @ (weight.float() if b_mn_major else weight.float().t())
            want = torch.empty(rpc, N, device=self.device, dtype=torch.float32)
            dist.reduce_scatter_tensor(want, full, group=self.group)
            err = (out.float() - want).abs().max().item()
            if not err <= 3e-2 * max(want.abs().max().item(), 1.0):
                raise RuntimeError(f"fused GEMM reduce-scatter (call {self.rs_calls}, rank {self.rank}): "
                                   f"max abs error {err}")
        return out

    # ---- GEMM -> all-reduce (tensor parallelism without sequence parallelism) ---------------
    def gemm_allreduce(self, a: torch.Tensor, weight: torch.Tensor, b_mn_major: bool = False):
        """a [M, K_local], weight [N, K_local] (or [K_local, N]); returns the [M, N] sum over ranks, the
        same bits on every rank (each row chunk is reduced once, by its owner, and stored to all)."""
        assert self.allreduce, "workspace was created without all-reduce output buffers"
        M, K = a.shape
        N = weight.shape[1] if b_mn_major else weight.shape[0]
        assert a.dtype == torch.bfloat16 and a.stride(1) == 1
        assert M <= self.max_rows and N <= self.max_n, (M, N, self.max_rows, self.max_n)
        self.rs_calls += 1
        s = self.rs_calls % 2
        parts, outs = (C.c_void_p * 8)(), (C.c_void_p * 8)()
        for r in range(self.tp):
            parts[r] = self._parts(r, s)
            outs[r] = self._out(r, s)
        L.call("b200_gemm_allreduce", C.byref(self._descs[0]), parts, outs, s, a.data_ptr(), weight.data_ptr(),
               M, N, K, int(b_mn_major), a.stride(0), weight.stride(0), L.stream_ptr())
        # the peers overwrite this buffer two calls from now: hand the caller its own copy
        out = _wrap(self._out(self.rank, s), M * N * 2, self.device).view(torch.bfloat16).view(M, N).clone()
        if self.verify:
            want = a.float() @ (weight.float() if b_mn_major else weight.float().t())
            dist.all_reduce(want, group=self.group)
            err = (out.float() - want).abs().max().item()
            if not err <= 3e-2 * max(want.abs().max().item(), 1.0):
                raise RuntimeError(f"fused GEMM all-reduce (call {self.rs_calls}, rank {self.rank}): max abs error {err}")
        return out

    # ---- staged all-gather window for kernels that pull rows themselves ----------------------
    def gather_begin(self, x_shard: torch.Tensor):
        """Make ``x_shard`` [rows, K] (same shape on every rank) readable by the peers and wait until the
        peers' shards are readable here. Returns (epoch, [address of rank r's rows]); the caller's
        kernel reads them and then calls ``gather_end(epoch)``. Shares the staging buffers and the
        epoch sequence of ``allgather_gemm``."""
        rows, K = x_shard.shape
        nbytes = rows * K * 2
        assert x_shard.is_contiguous() and x_shard.dtype == torch.bfloat16 and nbytes <= self.stage_bytes
        self.ag_epoch += 1
        self.stage_calls += 1
        e = self.ag_epoch
        stage = e % 2
        st = L.stream_ptr()
        addrs = (C.c_void_p * 8)()
        n = 0
        for r in range(self.tp):
            if r != self.rank:
                addrs[n] = self._flag(r, self.w_ready + self.rank)      # rank r's shard_ready[me]
                n += 1
        mine = self._stage(self.rank, stage)
        L.call("b200_stage_shard", x_shard.data_ptr(), mine, mine, nbytes, self._flag(self.rank, self.w_consumed),
               self.tp, self.rank, (e - 2) & 0xFFFFFFFF, int(e > 2), self._flag(self.rank, self.w_stage_done2),
               self.stage_calls, addrs, n, e, st)
        # the fused all-gather GEMM of the next epoch expects its counter set zeroed by this epoch
        nxt = (e & 1) ^ 1
        mb = self.max_blocks
        self._flags[nxt * mb:(nxt + 1) * mb].zero_()
        self._flags[2 * mb + 8 + nxt * 8:2 * mb + 16 + nxt * 8].zero_()
        L.call("b200_symm_wait", self._flag(self.rank, self.w_ready), self.tp, self.rank, e, st)
        srcs = [x_shard.data_ptr() if r == self.rank else self._stage(r, stage) for r in range(self.tp)]
        return e, srcs

    def gather_end(self, e: int):
        """Tell the peers this rank has finished reading their staging buffers of epoch ``e``."""
        addrs = (C.c_void_p * 8)()
        n = 0
        for r in range(self.tp):
            if r != self.rank:
                addrs[n] = self._flag(r, self.w_consumed + self.rank)   # rank r's consumed[me]
                n += 1
        L.call("b200_symm_signal", addrs, n, e, L.stream_ptr())

    # ---- vocab-parallel embedding tables ------------------------------------------------------
    def peer_tables(self, weight: torch.Tensor):
        """Addresses of every rank's shard of an embedding table in THIS rank's address space (CUDA IPC
        mapping of the allocation that holds the parameter; collective on first use per parameter).
        Returns None -- on EVERY rank -- when any rank could not export or map its allocation (memory
        from an allocator without IPC support, e.g. expandable segments): the caller then keeps the
        NCCL path. Both collectives below are reached by all ranks whatever happened locally."""
        key = (weight.data_ptr(), weight.numel())
        if key in self._tables:
            return self._tables[key]
        info = None
        try:
            handle = (C.c_ubyte * 64)()
            off = C.c_longlong()
            L.call("b200_ipc_export_range", weight.data_ptr(), handle, C.byref(off), launches=0)
            info = (bytes(handle), off.value)
        except Exception:  # noqa: BLE001 - reported to the group through `info is None`
            pass
        infos = [None] * self.tp
        dist.all_gather_object(infos, info, group=self.group)
        tabs = None
        if all(i is not None for i in infos):
            try:
                tabs = []
                for r, (h, o) in enumerate(infos):
                    if r == self.rank:
                        tabs.append(weight.data_ptr())
                        continue
                    base = _OPENED.get(h)
                    if base is None:
                        p = C.c_void_p()
                        L.call("b200_symm_import", (C.c_ubyte * 64).from_buffer_copy(h), C.byref(p), launches=0)
                        base = _OPENED[h] = p.value
                    tabs.append(base + o)
            except Exception:  # noqa: BLE001
                tabs = None
        oks = [None] * self.tp
        dist.all_gather_object(oks, tabs is not None, group=self.group)
        if not all(oks):
            tabs = None
        self._tables[key] = tabs
        return tabs

    def tables_barrier(self):
        """Stream-ordered handshake of the group: after it, every peer's kernels enqueued before ITS
        call (the optimizer step that wrote the table) are complete. One flag store per peer + one
        wait kernel."""
        self.table_calls += 1
        n_call = self.table_calls
        st = L.stream_ptr()
        addrs = (C.c_void_p * 8)()
        n = 0
        for r in range(self.tp):
            if r != self.rank:
                addrs[n] = self._flag(r, self.w_table_ready + self.rank)
                n += 1
        L.call("b200_symm_signal", addrs, n, n_call & 0xFFFFFFFF, st)
        L.call("b200_symm_wait", self._flag(self.rank, self.w_table_ready), self.tp, self.rank,
               n_call & 0xFFFFFFFF, st)

    # ---- vocab-parallel cross entropy: exchange of (max, sum-exp, target logit) ---------------
    def xent_exchange(self, logits2d: torch.Tensor, labels: torch.Tensor, vocab_start: int):
        """logits2d [T, V/tp] bf16, labels [T] int64 (global ids) -> (loss [T], lse [T]) fp32. The
        statistics kernel stores its triples into every rank's table and the combine kernel waits for
        all of them: two launches, no collective call."""
        T, V = logits2d.shape
        assert T <= self.max_rows, (T, self.max_rows)
        self.xent_calls += 1
        s = self.xent_calls % 2
        stats, cnts = (C.c_void_p * 8)(), (C.c_void_p * 8)()
        for r in range(self.tp):
            stats[r] = self._xent(r, s)
            cnts[r] = self._flag(r, self.w_xent_cnt + s)
        st = L.stream_ptr()
        L.call("b200_xent_stats_push", logits2d.data_ptr(), labels.data_ptr(), stats, cnts, self.tp, self.rank,
               T, V, vocab_start, st)
        self._xent_target[s] = (self._xent_target[s] + self.tp * L.lib().b200_xent_push_grid(T)) & 0xFFFFFFFF
        loss = torch.empty(T, device=self.device, dtype=torch.float32)
        lse = torch.empty(T, device=self.device, dtype=torch.float32)
        L.call("b200_xent_combine_wait", self._xent(self.rank, s), self._flag(self.rank, self.w_xent_cnt + s),
               self._xent_target[s], loss.data_ptr(), lse.data_ptr(), T, self.tp, st)
        if self.verify:
            # B200_COMM_VERIFY=1: the same statistics through ONE NCCL all-gather and the plain combine kernel
            stats = torch.empty(T, 3, device=self.device, dtype=torch.float32)
            L.call("b200_xent_stats", logits2d.data_ptr(), labels.data_ptr(), stats.data_ptr(), T, V, vocab_start, st)
            allst = torch.empty(self.tp * T, 3, device=self.device, dtype=torch.float32)
            dist.all_gather_into_tensor(allst, stats, group=self.group)
            want, want_lse = torch.empty_like(loss), torch.empty_like(lse)
            L.call("b200_xent_combine", allst.data_ptr(), want.data_ptr(), want_lse.data_ptr(), T, self.tp, st)
            if not (torch.allclose(loss, want, rtol=1e-5, atol=1e-5) and torch.allclose(lse, want_lse, rtol=1e-5, atol=1e-5)):
                raise RuntimeError(f"cross-entropy exchange over peer memory (call {self.xent_calls}, rank {self.rank}): "
                                   f"max loss difference {(loss - want).abs().max().item()} against NCCL")
        return loss, lse

    def check(self):
        """Synchronise and raise if a cross-GPU wait timed out since the last check (debug / tests)."""
        torch.cuda.synchronize()
        rec = (C.c_uint * 8)()
        L.call("b200_comm_debug_read", rec, launches=0)
        if rec[0]:
            site = {1: "row block not delivered by the copier warps", 2: "peer shard_ready flag",
                    3: "reduce-scatter arrivals"}.get(rec[0], str(rec[0]))
            raise RuntimeError(f"fused collective wait timed out: {site}; block {rec[1]}, saw {rec[2]}, "
                               f"wanted {rec[3]}, epoch {rec[4]}, mode {rec[5]}, rank {rec[6]}, index {rec[7]}")

    def close(self):
        torch.cuda.synchronize()
        for r, p in enumerate(self.peer_base):
            if r != self.rank:
                L.lib().b200_symm_unmap(C.c_void_p(p))
        L.lib().b200_symm_free(C.c_void_p(self.base))


_WORKSPACE = None


def get_workspace(group, max_rows, max_k, max_n, allreduce=False):
    """Process-wide workspace of the tensor-parallel group, grown on demand (collective: all
    ranks must ask for the same sizes)."""
    global _WORKSPACE
    ws = _WORKSPACE
    if ws is None or ws.max_rows < max_rows or ws.max_k < max_k or ws.max_n < max_n \
            or (allreduce and not ws.allreduce):
        if ws is not None:
            max_rows, max_k, max_n = max(max_rows, ws.max_rows), max(max_k, ws.max_k), max(max_n, ws.max_n)
            allreduce = allreduce or ws.allreduce
            # every GPU of the group must have drained its kernels (they read and write the peers'
            # buffers) BEFORE anybody unmaps or frees: device sync first, then the host barrier
            torch.cuda.synchronize()
            dist.barrier(group=group)
            ws.close()
        ws = _WORKSPACE = TPWorkspace(group, max_rows, max_k, max_n, allreduce)
    return ws
