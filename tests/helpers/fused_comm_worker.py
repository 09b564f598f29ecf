"""torchrun body: fused all-gather-GEMM / GEMM-reduce-scatter vs NCCL + matmul on the ranks of
one node. Prints RESULT json on rank 0."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist


def timed(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    from epfl_megatron_b200.parallel.symm import TPWorkspace
    grp = dist.group.WORLD
    bench = "--bench" in sys.argv
    shapes = [(512 * world, 1024, 768), (4096, 4096, 12288 // world), (4096, 4096, 22016 // world)] if world <= 4 \
        else [(256 * world, 1024, 768), (4096, 4096, 1536)]
    ws = TPWorkspace(grp, max_rows=4096, max_k=11008, max_n=4096, allreduce=True)
    out = {"world": world, "ag": [], "rs": [], "ar": []}
    torch.manual_seed(100 + rank)
    for (M, K, N) in shapes:
        rpc = M // world
        for b_mn in (False, True):
            x = (torch.randn(rpc, K, device="cuda") * 0.5).bfloat16()
            w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
            wst = w.t().contiguous() if b_mn else w
            errs = []
            for it in range(3):                      # repeated, no barriers in between
                y, gathered = ws.allgather_gemm(x, wst, b_mn)
                ws.check()
                full = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
                dist.all_gather_into_tensor(full, x)
                ref = full.float() @ w.float().t()
                assert torch.equal(gathered, full), "gathered rows differ"
                errs.append(float((y.float() - ref).abs().max() / ref.abs().max()))
                x = (x * 1.01).contiguous()
            rec = {"M": M, "K": K, "N": N, "b_mn": b_mn, "max_rel_err": max(errs)}
            if bench and not b_mn:
                def fused():
                    ws.allgather_gemm(x, wst, b_mn)

                def nccl():
                    full = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
                    dist.all_gather_into_tensor(full, x)
                    torch.matmul(full, w.t())
                rec["ms_fused"], rec["ms_nccl_cublas"] = timed(fused), timed(nccl)
                rec["ms_gemm_only"] = timed(lambda: torch.matmul(torch.empty(M, K, device="cuda", dtype=torch.bfloat16), w.t()))
            out["ag"].append(rec)
    # reduce-scatter: a [M, K/tp] per rank, w [N, K/tp]
    rs_shapes = [(512 * world, 1024, 1024), (4096, 4096 // world * 1, 4096), (4096, 11008 // world, 4096)] if world <= 4 \
        else [(256 * world, 512, 1024), (4096, 1408, 4096)]
    for (M, K, N) in rs_shapes:
        K = K // 8 * 8
        rpc = M // world
        for b_mn in (False, True):
            a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
            w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
            wst = w.t().contiguous() if b_mn else w
            errs = []
            for it in range(3):
                y = ws.gemm_reduce_scatter(a, wst, b_mn)
                ws.check()
                part = a.float() @ w.float().t()
                ref = torch.empty(rpc, N, device="cuda")
                dist.reduce_scatter_tensor(ref, part)
                errs.append(float((y.float() - ref).abs().max() / ref.abs().max()))
                a = (a * 0.99).contiguous()
            rec = {"M": M, "K": K, "N": N, "b_mn": b_mn, "max_rel_err": max(errs)}
            if bench and not b_mn:
                def fused():
                    ws.gemm_reduce_scatter(a, wst, b_mn)

                def nccl():
                    p = torch.matmul(a, w.t())
                    o = torch.empty(rpc, N, device="cuda", dtype=torch.bfloat16)
                    dist.reduce_scatter_tensor(o, p)
                rec["ms_fused"], rec["ms_nccl_cublas"] = timed(fused), timed(nccl)
                rec["ms_gemm_only"] = timed(lambda: torch.matmul(a, w.t()))
            out["rs"].append(rec)
    # all-reduce (tensor parallelism without sequence parallelism): same shapes + one ragged width
    for (M, K, N) in rs_shapes + [(256 * world, 520, 1000)]:
        K = K // 8 * 8
        for b_mn in (False, True):
            a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
            w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
            wst = w.t().contiguous() if b_mn else w
            errs = []
            for it in range(3):
                y = ws.gemm_allreduce(a, wst, b_mn)
                ws.check()
                ref = a.float() @ w.float().t()
                dist.all_reduce(ref)
                errs.append(float((y.float() - ref).abs().max() / ref.abs().max()))
                ys = [torch.empty_like(y) for _ in range(world)]
                dist.all_gather(ys, y)
                assert all(torch.equal(ys[0], t) for t in ys), "all-reduce GEMM: ranks hold different bits"
                a = (a * 0.99).contiguous()
            rec = {"M": M, "K": K, "N": N, "b_mn": b_mn, "max_rel_err": max(errs)}
            if bench and not b_mn and N % 256 == 0:
                def fused():
                    ws.gemm_allreduce(a, wst, b_mn)

                def nccl():
                    p = torch.matmul(a, w.t())
                    dist.all_reduce(p)
                rec["ms_fused"], rec["ms_nccl_cublas"] = timed(fused), timed(nccl)
                rec["ms_gemm_only"] = timed(lambda: torch.matmul(a, w.t()))
            out["ar"].append(rec)
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print("RESULT " + json.dumps(out), flush=True)
    ws.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
