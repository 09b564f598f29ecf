"""Multi-process CPU harness: run a top-level function on ``world`` Gloo ranks."""
import os
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from .launch import free_port


def _worker(rank, world, port, fn, args, tp, pp, vpp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.manual_seed(1234)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from epfl_megatron_b200.parallel import state as mpu
        from epfl_megatron_b200.parallel.random import model_parallel_cuda_manual_seed
        mpu.destroy_model_parallel()
        mpu.initialize_model_parallel(tp, pp, vpp)
        model_parallel_cuda_manual_seed(1234)
        fn(rank, world, *args)
    except Exception:
        traceback.print_exc()
        raise
    finally:
        dist.barrier()
        dist.destroy_process_group()


def spawn(fn, world, *args, tp=1, pp=1, vpp=None):
    mp.spawn(_worker, args=(world, free_port(), fn, args, tp, pp, vpp), nprocs=world, join=True)
