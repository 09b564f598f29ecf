"""Process-group topology: tensor (fastest), data, pipeline (slowest) axes.

Same rank layout and getter/setter surface as megatron/core/parallel_state.py:51-524
(aliased ``mpu`` in the reference): TP groups are contiguous ranks, DP groups stride by
tp inside a pipeline block, PP groups stride by world/pp; the embedding group is
{first, last[, split]} pipeline rank. Works on NCCL and Gloo. On one 8xB200 box every
group is inside a single NVSwitch domain, so no hierarchy is needed.
"""
from __future__ import annotations

from typing import List, Optional

import os

import torch
import torch.distributed as dist

_TP = _PP = _DP = _MP = _EMB = _POS_EMB = None
_TP_CPU = None
_TP_RANKS: Optional[List[int]] = None
_PP_RANKS: Optional[List[int]] = None
_DP_RANKS: Optional[List[int]] = None
_EMB_RANKS: Optional[List[int]] = None
_POS_EMB_RANKS: Optional[List[int]] = None

_VPP_RANK = None
_VPP_WORLD = None
_PP_SPLIT_RANK = None

# overrides used by offline tools that fake ranks in one process
_OVR = {"tp_ws": None, "pp_ws": None, "tp_rank": None, "pp_rank": None, "dp_ws": None, "dp_rank": None}

_GLOBAL_MEMORY_BUFFER = None


def rank_layout(world_size: int, tp: int, pp: int):
    """Pure function: lists of rank lists for every kind of group (no communication)."""
    if world_size % (tp * pp) != 0:
        raise RuntimeError(f"world_size ({world_size}) is not divisible by tp ({tp}) x pp ({pp})")
    dp = world_size // (tp * pp)
    per_stage = world_size // pp
    tp_groups = [list(range(i * tp, (i + 1) * tp)) for i in range(world_size // tp)]
    dp_groups = []
    for stage in range(pp):
        lo = stage * per_stage
        for t in range(tp):
            dp_groups.append(list(range(lo + t, lo + per_stage, tp)))
    mp_groups = []
    for d in range(dp):
        mp_groups.append(sorted(g[d] for g in dp_groups))
    pp_groups = [list(range(i, world_size, per_stage)) for i in range(per_stage)]
    return dict(tp=tp_groups, dp=dp_groups, mp=mp_groups, pp=pp_groups, dp_size=dp)


def initialize_model_parallel(tensor_model_parallel_size: int = 1,
                              pipeline_model_parallel_size: int = 1,
                              virtual_pipeline_model_parallel_size: Optional[int] = None,
                              pipeline_model_parallel_split_rank: Optional[int] = None) -> None:
    global _TP, _PP, _DP, _MP, _EMB, _POS_EMB, _TP_RANKS, _PP_RANKS, _DP_RANKS
    global _EMB_RANKS, _POS_EMB_RANKS, _VPP_RANK, _VPP_WORLD, _PP_SPLIT_RANK
    assert dist.is_initialized(), "torch.distributed is not initialized"
    assert _TP is None, "model parallel groups are already initialized"
    world, rank = dist.get_world_size(), dist.get_rank()
    lay = rank_layout(world, tensor_model_parallel_size, pipeline_model_parallel_size)
    if virtual_pipeline_model_parallel_size is not None:
        if pipeline_model_parallel_size < 2:
            raise RuntimeError("interleaved schedule needs pipeline_model_parallel_size >= 2")
        _VPP_RANK, _VPP_WORLD = 0, virtual_pipeline_model_parallel_size
    _PP_SPLIT_RANK = pipeline_model_parallel_split_rank

    def make(groups):
        mine, mine_ranks = None, None
        for ranks in groups:  # every rank must create every group, in the same order
            g = dist.new_group(ranks)
            if rank in ranks:
                mine, mine_ranks = g, ranks
        return mine, mine_ranks

    _DP, _DP_RANKS = make(lay["dp"])
    _MP, _ = make(lay["mp"])
    _TP, _TP_RANKS = make(lay["tp"])
    # Control plane of the tensor-parallel group on the HOST (Gloo): the per-micro-batch shape
    # exchange of broadcast_data goes through it, so that it never waits for the GPU stream.
    global _TP_CPU
    _TP_CPU = None
    if tensor_model_parallel_size > 1 and dist.get_backend() == "nccl" \
            and os.environ.get("B200_TP_CPU_GROUP", "1") != "0":
        try:
            for ranks in lay["tp"]:
                g = dist.new_group(ranks, backend="gloo")
                if rank in ranks:
                    _TP_CPU = g
        except Exception as e:  # noqa: BLE001 - Gloo cannot come up here (no usable interface): NCCL path
            print(f"[rank {rank}] no host-side tensor-parallel group ({type(e).__name__}: {e}); "
                  "broadcast_data falls back to NCCL for the shapes", flush=True)
            _TP_CPU = None
    _PP, _PP_RANKS = make(lay["pp"])
    emb_groups, pos_groups = [], []
    for ranks in lay["pp"]:
        if len(ranks) > 1:
            e, p = [ranks[0], ranks[-1]], [ranks[0]]
            s = pipeline_model_parallel_split_rank
            if s is not None:
                if ranks[s] not in e:
                    e = [ranks[0], ranks[s], ranks[-1]]
                if ranks[s] not in p:
                    p = [ranks[0], ranks[s]]
        else:
            e, p = list(ranks), list(ranks)
        emb_groups.append(e)
        pos_groups.append(p)
    _EMB, _EMB_RANKS = make(emb_groups)
    _POS_EMB, _POS_EMB_RANKS = make(pos_groups)
    # a middle pipeline stage belongs to neither group: an empty list, not the "uninitialised" None
    _EMB_RANKS = _EMB_RANKS or []
    _POS_EMB_RANKS = _POS_EMB_RANKS or []
    _set_global_memory_buffer()


def get_tensor_model_parallel_cpu_group():
    """Gloo group over the ranks of the tensor-parallel group, or None (single rank, Gloo world, disabled)."""
    return _TP_CPU


def model_parallel_is_initialized() -> bool:
    return _TP is not None and _PP is not None and _DP is not None


def _need(g, what):
    assert g is not None, f"{what} group is not initialized"
    return g


def get_model_parallel_group():
    return _need(_MP, "model parallel")


def get_tensor_model_parallel_group():
    return _need(_TP, "tensor model parallel")


def get_pipeline_model_parallel_group():
    return _need(_PP, "pipeline model parallel")


def get_data_parallel_group():
    return _need(_DP, "data parallel")


def get_embedding_group():
    return _need(_EMB, "embedding")


def get_position_embedding_group():
    return _need(_POS_EMB, "position embedding")


def get_tensor_model_parallel_ranks():
    return _TP_RANKS


def get_pipeline_model_parallel_ranks():
    return _PP_RANKS


def get_data_parallel_ranks():
    return _DP_RANKS


# ---- overrides (tools/checkpoint_util fakes ranks in a single CPU process) ----
def set_tensor_model_parallel_world_size(n):
    _OVR["tp_ws"] = n


def set_pipeline_model_parallel_world_size(n):
    _OVR["pp_ws"] = n


def set_tensor_model_parallel_rank(r):
    _OVR["tp_rank"] = r


def set_pipeline_model_parallel_rank(r):
    _OVR["pp_rank"] = r


def set_data_parallel_world_size(n):
    _OVR["dp_ws"] = n


def set_data_parallel_rank(r):
    _OVR["dp_rank"] = r


def get_tensor_model_parallel_world_size():
    if _OVR["tp_ws"] is not None:
        return _OVR["tp_ws"]
    if _TP is None and not dist.is_initialized():
        return 1
    return dist.get_world_size(group=get_tensor_model_parallel_group())


def get_pipeline_model_parallel_world_size():
    if _OVR["pp_ws"] is not None:
        return _OVR["pp_ws"]
    if _PP is None and not dist.is_initialized():
        return 1
    return dist.get_world_size(group=get_pipeline_model_parallel_group())


def get_tensor_model_parallel_rank():
    if _OVR["tp_rank"] is not None:
        return _OVR["tp_rank"]
    if _TP is None and not dist.is_initialized():
        return 0
    return dist.get_rank(group=get_tensor_model_parallel_group())


def get_pipeline_model_parallel_rank():
    if _OVR["pp_rank"] is not None:
        return _OVR["pp_rank"]
    if _PP is None and not dist.is_initialized():
        return 0
    return dist.get_rank(group=get_pipeline_model_parallel_group())


def get_data_parallel_world_size():
    if _OVR["dp_ws"] is not None:
        return _OVR["dp_ws"]
    if _DP is None and not dist.is_initialized():
        return 1
    return dist.get_world_size(group=get_data_parallel_group())


def get_data_parallel_rank():
    if _OVR["dp_rank"] is not None:
        return _OVR["dp_rank"]
    if _DP is None and not dist.is_initialized():
        return 0
    return dist.get_rank(group=get_data_parallel_group())


def is_pipeline_first_stage(ignore_virtual: bool = False) -> bool:
    if not ignore_virtual and _VPP_WORLD is not None and _VPP_RANK != 0:
        return False
    return get_pipeline_model_parallel_rank() == 0


def is_pipeline_last_stage(ignore_virtual: bool = False) -> bool:
    if not ignore_virtual and _VPP_WORLD is not None and _VPP_RANK != _VPP_WORLD - 1:
        return False
    return get_pipeline_model_parallel_rank() == get_pipeline_model_parallel_world_size() - 1


def is_rank_in_embedding_group(ignore_virtual: bool = False) -> bool:
    rank = dist.get_rank() if dist.is_initialized() else 0
    if _EMB_RANKS is None:
        return True
    if ignore_virtual:
        return rank in _EMB_RANKS
    if rank in _EMB_RANKS:
        if rank == _EMB_RANKS[0]:
            return is_pipeline_first_stage()
        if rank == _EMB_RANKS[-1]:
            return is_pipeline_last_stage()
        return True
    return False


def is_rank_in_position_embedding_group() -> bool:
    rank = dist.get_rank() if dist.is_initialized() else 0
    return _POS_EMB_RANKS is None or rank in _POS_EMB_RANKS


def get_pipeline_model_parallel_split_rank():
    return _PP_SPLIT_RANK


def set_pipeline_model_parallel_split_rank(r):
    global _PP_SPLIT_RANK
    _PP_SPLIT_RANK = r


def is_pipeline_stage_before_split(rank=None) -> bool:
    if get_pipeline_model_parallel_world_size() == 1 or _PP_SPLIT_RANK is None:
        return True
    rank = get_pipeline_model_parallel_rank() if rank is None else rank
    return rank < _PP_SPLIT_RANK


def is_pipeline_stage_after_split(rank=None) -> bool:
    if get_pipeline_model_parallel_world_size() == 1 or _PP_SPLIT_RANK is None:
        return True
    rank = get_pipeline_model_parallel_rank() if rank is None else rank
    return rank >= _PP_SPLIT_RANK


def is_pipeline_stage_at_split() -> bool:
    r = get_pipeline_model_parallel_rank()
    return is_pipeline_stage_before_split(r) and is_pipeline_stage_after_split(r + 1)


def get_virtual_pipeline_model_parallel_rank():
    return _VPP_RANK


def set_virtual_pipeline_model_parallel_rank(r):
    global _VPP_RANK
    _VPP_RANK = r


def get_virtual_pipeline_model_parallel_world_size():
    return _VPP_WORLD


def set_virtual_pipeline_model_parallel_world_size(n):
    global _VPP_WORLD
    _VPP_WORLD = n


def get_tensor_model_parallel_src_rank():
    rank = dist.get_rank() if dist.is_initialized() else 0
    tp = get_tensor_model_parallel_world_size()
    return (rank // tp) * tp


def get_data_parallel_src_rank():
    return _DP_RANKS[0] if _DP_RANKS else 0


def get_pipeline_model_parallel_first_rank():
    return _PP_RANKS[0]


def get_pipeline_model_parallel_last_rank():
    return _PP_RANKS[get_pipeline_model_parallel_world_size() - 1]


def get_pipeline_model_parallel_next_rank():
    r, n = get_pipeline_model_parallel_rank(), get_pipeline_model_parallel_world_size()
    return _PP_RANKS[(r + 1) % n]


def get_pipeline_model_parallel_prev_rank():
    r, n = get_pipeline_model_parallel_rank(), get_pipeline_model_parallel_world_size()
    return _PP_RANKS[(r - 1) % n]


class GlobalMemoryBuffer:
    """Grow-only named scratch tensors (megatron/core/utils.py:24-45 behaviour)."""

    def __init__(self):
        self._buf = {}

    def get_tensor(self, shape, dtype, name):
        need = 1
        for s in shape:
            need *= int(s)
        key = (name, dtype)
        cur = self._buf.get(key)
        if cur is None or cur.numel() < need:
            dev = torch.cuda.current_device() if torch.cuda.is_available() else "cpu"
            cur = torch.empty(need, dtype=dtype, device=dev, requires_grad=False)
            self._buf[key] = cur
        return cur[:need].view(*shape)


def _set_global_memory_buffer():
    global _GLOBAL_MEMORY_BUFFER
    _GLOBAL_MEMORY_BUFFER = GlobalMemoryBuffer()


def get_global_memory_buffer():
    global _GLOBAL_MEMORY_BUFFER
    if _GLOBAL_MEMORY_BUFFER is None:
        _set_global_memory_buffer()
    return _GLOBAL_MEMORY_BUFFER


def destroy_model_parallel():
    global _TP, _PP, _DP, _MP, _EMB, _POS_EMB, _TP_RANKS, _PP_RANKS, _DP_RANKS
    global _EMB_RANKS, _POS_EMB_RANKS, _VPP_RANK, _VPP_WORLD, _PP_SPLIT_RANK, _GLOBAL_MEMORY_BUFFER
    global _TP_CPU
    _TP_CPU = None
    _TP = _PP = _DP = _MP = _EMB = _POS_EMB = None
    _TP_RANKS = _PP_RANKS = _DP_RANKS = _EMB_RANKS = _POS_EMB_RANKS = None
    _VPP_RANK = _VPP_WORLD = _PP_SPLIT_RANK = None
    _GLOBAL_MEMORY_BUFFER = None
    for k in _OVR:
        _OVR[k] = None
