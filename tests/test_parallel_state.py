"""Group construction and getters (ideas of the reference's tests/test_parallel_state.py,
made world-size parametric and runnable on CPU/Gloo)."""
import pytest
import torch.distributed as dist

from epfl_megatron_b200.parallel import state as mpu
from tests.helpers.dist import spawn


@pytest.mark.parametrize("world,tp,pp", [(8, 4, 2), (8, 2, 4), (8, 2, 1), (16, 2, 4), (4, 1, 1), (8, 8, 1)])
def test_rank_layout_math(world, tp, pp):
    lay = mpu.rank_layout(world, tp, pp)
    dp = world // (tp * pp)
    assert lay["dp_size"] == dp
    assert all(g == list(range(g[0], g[0] + tp)) for g in lay["tp"])          # TP ranks adjacent
    assert all(len(g) == dp and all(b - a == tp for a, b in zip(g, g[1:])) for g in lay["dp"])
    assert all(len(g) == pp and all(b - a == world // pp for a, b in zip(g, g[1:])) for g in lay["pp"])
    for kind in ("tp", "dp", "pp"):
        flat = sorted(r for g in lay[kind] for r in g)
        assert flat == list(range(world))                                       # a partition
    assert all(len(g) == tp * pp for g in lay["mp"])


def test_docstring_example_16_ranks():
    lay = mpu.rank_layout(16, 2, 4)
    assert lay["tp"][0] == [0, 1] and lay["tp"][7] == [14, 15]
    assert [0, 2] in lay["dp"] and [13, 15] in lay["dp"]
    assert [0, 4, 8, 12] in lay["pp"] and [3, 7, 11, 15] in lay["pp"]


def test_indivisible_world_raises():
    with pytest.raises(RuntimeError):
        mpu.rank_layout(8, 3, 1)


def _check_groups(rank, world, tp, pp):
    assert mpu.model_parallel_is_initialized()
    assert mpu.get_tensor_model_parallel_world_size() == tp
    assert mpu.get_pipeline_model_parallel_world_size() == pp
    assert mpu.get_data_parallel_world_size() == world // (tp * pp)
    assert mpu.get_tensor_model_parallel_rank() == rank % tp
    assert mpu.get_pipeline_model_parallel_rank() == rank // (world // pp)
    assert mpu.get_tensor_model_parallel_src_rank() == (rank // tp) * tp
    assert mpu.is_pipeline_first_stage() == (rank // (world // pp) == 0)
    assert mpu.is_pipeline_last_stage() == (rank // (world // pp) == pp - 1)
    if pp > 1:
        assert mpu.get_pipeline_model_parallel_next_rank() == (rank + world // pp) % world
        assert mpu.get_pipeline_model_parallel_prev_rank() == (rank - world // pp) % world
    # overrides used by the offline tools
    mpu.set_tensor_model_parallel_world_size(7)
    mpu.set_tensor_model_parallel_rank(3)
    assert mpu.get_tensor_model_parallel_world_size() == 7 and mpu.get_tensor_model_parallel_rank() == 3
    mpu.set_tensor_model_parallel_world_size(None)
    mpu.set_tensor_model_parallel_rank(None)
    mpu.set_virtual_pipeline_model_parallel_rank(1)
    assert mpu.get_virtual_pipeline_model_parallel_rank() == 1
    mpu.set_virtual_pipeline_model_parallel_rank(None)
    mpu.destroy_model_parallel()
    assert not mpu.model_parallel_is_initialized()


@pytest.mark.parametrize("world,tp,pp", [(4, 2, 2), (2, 2, 1), (4, 1, 4)])
def test_groups_on_gloo(world, tp, pp):
    spawn(_check_groups, world, tp, pp, tp=tp, pp=pp)
