"""Headline benchmark: Llama-2 7B bf16 training throughput (tokens/s, whole job).

    python bench.py --gpus N --steps K --warmup W            (N = 1)
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Config per BASELINE.md section 4: N=1 -> TP1, N=2 -> TP2, N=4 -> TP2xDP2, N=8 -> TP2xDP4;
seq 4096, micro batch 1, 4 micro-batches per step and DP rank (weak scaling over DP),
sequence parallel, distributed optimizer, fused attention, synthetic tokens, random init.
The step goes through the public API (``initialize_megatron`` ->
``setup_model_and_optimizer`` -> ``train_step``): forward, backward, grad reduction,
optimizer step, param gather. Inputs come from pinned host memory every step.

Prints ONE JSON line on rank 0. ``--impl reference`` runs the unmodified reference from
``baseline/_ref`` on the same config (see baseline/run_reference.py).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

MODELS = {
    # name: layers, hidden, heads, kv heads, ffn, vocab
    "llama2-7b": (32, 4096, 32, 32, 11008, 32000),
    "llama2-13b": (40, 5120, 40, 40, 13824, 32000),
    "llama2-70b": (80, 8192, 64, 8, 28672, 32000),
    "tiny": (2, 256, 8, 2, 688, 32000),
    # Falcon: LayerNorm, GeLU MLP of 4h, parallel attention / MLP block, tied embeddings
    "falcon-40b": (60, 8192, 128, 8, 32768, 65024),
}


# tokens/s of the UNMODIFIED reference (``--impl reference``) on one 8 x B200 box with the default
# layout of this file at N GPUs; the denominator of ``vs_baseline``.
REF_SAME_BOX = {1: 13.6e3, 2: 21.7e3, 4: 43.2e3, 8: 85.6e3}


def parallel_layout(n_gpus, model, tp_override=0, pp=1):
    if tp_override:
        return tp_override, n_gpus // (tp_override * pp)
    if model == "tiny":
        return (1, n_gpus) if n_gpus < 2 else (2, n_gpus // 2)
    tp = 1 if n_gpus == 1 else 2
    return tp, n_gpus // tp


def build_arg_list(a, tp, dp):
    L, h, nh, nkv, ffn, vocab = MODELS[a.model]
    if getattr(a, 'debug_layers', 0):
        L = a.debug_layers
    gbs = a.micro_batch * a.num_microbatches * dp
    argv = [
        "--b200_kernels", getattr(a, 'kernels', 'auto'), "--b200_fused_comm", getattr(a, 'fused_comm', 'auto'),
        "--b200_fused_dp", getattr(a, 'fused_dp', 'auto'), "--ddp_bucket_mb", str(getattr(a, 'bucket_mb', 256)),
        "--model_name", "falcon" if a.model.startswith("falcon") else "llama2",
        "--num_layers", str(L), "--hidden_size", str(h),
        "--num_attention_heads", str(nh), "--num_attention_heads_kv", str(nkv),
        "--ffn_hidden_size", str(ffn), "--seq_length", str(a.seq), "--max_position_embeddings", str(a.seq),
        "--micro_batch_size", str(a.micro_batch), "--global_batch_size", str(gbs),
        "--train_iters", str(a.steps * 2 + a.warmup + 8), "--lr", "1e-5", "--min_lr", "1e-6",
        "--lr_decay_style", "cosine", "--lr_warmup_iters", "1", "--weight_decay", "0.1",
        "--adam_beta2", "0.95", "--clip_grad", "1.0",
        *(["--parallel_attn", "--parallel_layernorm"] if a.model.startswith("falcon") else
          ["--use_rms_norm", "--glu_activation", "swiglu", "--no_tie_embed_logits"]),
        "--position_embedding_type", "rotary", "--use_flash_attn", "--no_bias_gelu_fusion",
        "--hidden_dropout", "0.0", "--attention_dropout", "0.0", "--bf16",
        "--tensor_model_parallel_size", str(tp), "--pipeline_model_parallel_size", str(getattr(a, 'pp', 1)),
        "--use_distributed_optimizer", "--tokenizer_type", "NullTokenizer",
        "--make_vocab_size_divisible_by", "128", "--synthetic_data", "--num_workers", "0",
        "--log_interval", "1000000", "--eval_iters", "0", "--eval_interval", "1000000",
    ]
    if tp > 1 and not getattr(a, 'no_sp', False):
        argv.append("--sequence_parallel")
    if getattr(a, 'fp8', False):
        argv += ["--fp8_hybrid", "--fp8_amax_history_len", "16", "--fp8_amax_compute_algo", "max"]
        argv += os.environ.get("B200_BENCH_FP8_EXTRA", "").split()
    if a.recompute:
        argv += ["--recompute_granularity", "full", "--recompute_method", "uniform"]
    if getattr(a, 'vpp_layers', 0):
        argv += ["--num_layers_per_virtual_pipeline_stage", str(a.vpp_layers)]
    return argv, gbs


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""
    REASONS = {0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x8: "hw_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            while not self._halt.is_set():
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                bits = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
                time.sleep(0.2)
        except Exception as e:  # noqa: BLE001
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def run_ours(a):
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    import finetune
    from epfl_megatron_b200 import get_args, get_timers
    from epfl_megatron_b200.initialize import initialize_megatron
    from epfl_megatron_b200.models.enums import ModelType
    from epfl_megatron_b200.ops import lib as klib
    from epfl_megatron_b200.parallel import state as mpu
    from epfl_megatron_b200.training import model_flops_per_token, setup_model_and_optimizer, train_step

    world = int(os.environ.get("WORLD_SIZE", "1"))
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}"
    tp, dp = parallel_layout(a.gpus, a.model, getattr(a, 'tp', 0), getattr(a, 'pp', 1))
    argv, gbs = build_arg_list(a, tp, dp)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29400")
    quiet = os.environ.get("RANK", "0") != "0" or not a.verbose
    if quiet:
        devnull = open(os.devnull, "w")
        real_stdout, sys.stdout = sys.stdout, devnull
    initialize_megatron(finetune.extra_args, {}, args_list=argv)
    args = get_args()
    model, optimizer, sched = setup_model_and_optimizer(finetune.model_provider, ModelType.encoder_or_decoder)
    for m in model:
        m.train()
    args.do_train = True
    dev = torch.device("cuda", torch.cuda.current_device())
    rank = dist.get_rank()

    # synthetic batches in pinned host memory; every step copies its inputs host -> device
    n_mb = args.global_batch_size // (args.micro_batch_size * dp)
    gen = torch.Generator().manual_seed(1234 + mpu.get_data_parallel_rank())
    pool = [torch.randint(0, args.padded_vocab_size, (args.micro_batch_size, args.seq_length + 1),
                          generator=gen, dtype=torch.int64).pin_memory() for _ in range(8)]

    def batches():
        i = 0
        while True:
            yield {"text": pool[i % len(pool)]}
            i += 1

    it = batches() if mpu.get_tensor_model_parallel_rank() == 0 else None
    if args.virtual_pipeline_model_parallel_size is not None:
        it = [batches() if mpu.get_tensor_model_parallel_rank() == 0 else None for _ in range(len(model))]
    h2d_bytes = n_mb * pool[0].numel() * pool[0].element_size()

    def step():
        return train_step(finetune.forward_step, it, model, optimizer, sched, args)

    dbg = []
    for _ in range(a.warmup):
        r = step()
        if os.environ.get('B200_BENCH_DEBUG_LOSSES') and r[0]:
            dbg.append(float(r[0]['lm loss'].item()))
    if dbg and rank == 0:
        print('DEBUG warmup losses', dbg, file=sys.stderr, flush=True)
    torch.cuda.synchronize()
    dist.barrier()

    # ---- region A: device-timed, no host read-back ----
    sampler = ClockSampler(torch.cuda.current_device())
    sampler.start()
    launches0 = klib.launch_count()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = e0.elapsed_time(e1)
    launches = klib.launch_count() - launches0

    if a.torch_profile:
        # one step under the torch profiler: per-kernel device time and the GPU-busy fraction
        from torch.profiler import ProfilerActivity, profile
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            step()
            torch.cuda.synchronize()
        if rank == 0:
            ev = [e for e in prof.key_averages() if e.device_time_total > 0]
            ev.sort(key=lambda e: -e.device_time_total)
            tot = sum(e.device_time_total for e in ev)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            with open(os.path.join(ROOT, "gpurun_out", f"torch_profile_n{a.gpus}.md"), "w") as f:
                f.write(f"# one training step, {a.model}, {a.gpus} GPU(s): device time by kernel\n\n")
                f.write(f"step wall (device-timed, outside the profiler): {ms / a.steps:.1f} ms; "
                        f"sum of kernel time under the profiler: {tot / 1e3:.1f} ms\n\n")
                f.write("| share | total ms | calls | avg us | kernel |\n|---|---|---|---|---|\n")
                for e in ev[:60]:
                    f.write(f"| {e.device_time_total / tot * 100:.2f}% | {e.device_time_total / 1e3:.2f} | {e.count} | "
                            f"{e.device_time_total / max(e.count, 1):.1f} | `{e.key[:110]}` |\n")
                # GPU-busy fraction: union of the kernel intervals of this rank, and where the gaps are
                try:
                    from torch.autograd import DeviceType
                    ks = sorted(((x.time_range.start, x.time_range.end, x.name) for x in prof.events()
                                 if x.device_type == DeviceType.CUDA), key=lambda t: t[0])
                    busy, gaps, cur_end, prev = 0.0, {}, None, None
                    for st_, en_, nm in ks:
                        if cur_end is None:
                            cur_end = st_
                        if st_ > cur_end:
                            g = gaps.setdefault(prev, [0, 0.0])
                            g[0] += 1
                            g[1] += st_ - cur_end
                            cur_end = st_
                        if en_ > cur_end:
                            busy += en_ - cur_end
                            cur_end = en_
                        prev = nm[:70]
                    span = ks[-1][1] - ks[0][0]
                    f.write(f"\nGPU busy {busy / 1e3:.1f} ms of a {span / 1e3:.1f} ms span under the profiler "
                            f"({busy / span * 100:.1f} %); idle {(span - busy) / 1e3:.1f} ms. Idle time by the kernel "
                            "that ran BEFORE the gap:\n\n| idle ms | gaps | avg us | preceding kernel |\n|---|---|---|---|\n")
                    for nm, (cnt, tot_g) in sorted(gaps.items(), key=lambda kv: -kv[1][1])[:25]:
                        f.write(f"| {tot_g / 1e3:.2f} | {cnt} | {tot_g / cnt:.1f} | `{nm}` |\n")
                except Exception as ex:  # noqa: BLE001
                    f.write(f"\n(gap analysis failed: {ex})\n")
    # ---- region B: end to end (pinned H2D inputs + D2H loss every step), wall clock ----
    if a.profile:
        sampler.stop()
        if rank == 0:
            sys.stdout = real_stdout if quiet else sys.stdout
            print(json.dumps({"profile_only": True, "ms_per_step": ms / a.steps}))
        dist.barrier()
        dist.destroy_process_group()
        return
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    # Every step copies its loss device -> pinned host memory (4 bytes) and the host reads it; like a
    # training loop that logs, the read of step i happens after step i+1 has been enqueued, so the
    # launch queue never runs dry. All reads are inside the timed region.
    last_loss, losses = None, []
    host_loss = torch.zeros(a.steps, dtype=torch.float32).pin_memory()
    done = [torch.cuda.Event() for _ in range(a.steps)]
    have = [False] * a.steps

    def read(i):
        done[i].synchronize()
        if have[i]:
            losses.append(float(host_loss[i]))
    for i in range(a.steps):
        loss_dict, _, _, _ = step()
        if loss_dict:
            host_loss[i:i + 1].copy_(loss_dict["lm loss"].detach().float().reshape(1), non_blocking=True)
            have[i] = True
        done[i].record()
        if i > 0:
            read(i - 1)
    read(a.steps - 1)
    last_loss = losses[-1] if losses else None
    torch.cuda.synchronize()
    dist.barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()

    t = torch.tensor([ms, e2e_s * 1000.0], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(t[0]), float(t[1])
    tokens_per_step = args.global_batch_size * args.seq_length
    value = tokens_per_step * a.steps / (ms_max / 1000.0)
    e2e_value = tokens_per_step * a.steps / (e2e_ms_max / 1000.0)
    flops_tok = model_flops_per_token(args)
    peaks = {"bf16_tflops_sustained": 1400.0, "bf16_tflops": 1590.0}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks.update(json.load(f))
    except OSError:
        pass
    tflops_gpu = value * flops_tok / 1e12 / a.gpus
    if quiet:
        sys.stdout = real_stdout
    if rank == 0:
        custom_layout = bool(getattr(a, 'fp8', False) or getattr(a, 'no_sp', False) or getattr(a, 'tp', 0) or getattr(a, 'pp', 1) > 1 or a.seq != 4096 or a.recompute)
        out = {
            "metric": ("Llama-2 7B" if a.model == "llama2-7b" else a.model) +
                      " bf16 training throughput (tokens/s, whole job, max over ranks)",
            "value": value, "unit": "tokens/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / REF_SAME_BOX[a.gpus] if a.model == "llama2-7b" and not custom_layout
                            and REF_SAME_BOX.get(a.gpus) else None),
            "vs_published_a100_anchor": value / 7100.0 if a.model == "llama2-7b" else None,
            "dtype": "fp8 (hybrid, delayed scaling) linears on bf16 parameters" if getattr(a, 'fp8', False) else "bf16",
            "data": "synthetic tokens, random-init weights",
            "impl": "ours",
            "config": {"model": a.model, "global_batch": args.global_batch_size, "seq_len": args.seq_length,
                       "micro_batch": args.micro_batch_size, "num_microbatches": n_mb,
                       "parallelism": f"tp{tp}" + (f"pp{a.pp}" if a.pp > 1 else "") + (f"dp{dp}" if dp > 1 else "")
                                      + (f" interleaved x{len(model)}" if len(model) > 1 else "")
                                      + (" full-recompute" if a.recompute else ""),
                       "sequence_parallel": bool(args.sequence_parallel), "distributed_optimizer": True,
                       "l2_policy": "inputs larger than L2 (13.5 GB of weights streamed every step)",
                       "baseline_note": "vs_baseline divides by the reference's own tokens/s measured on a B200 box "
                                        "with this config (BASELINE.md section 4); vs_published_a100_anchor by "
                                        "the 7.1k tokens/s derived from its tutorial (8xA100, seq 1024)"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "tokens/s", "ms_per_step": e2e_ms_max / a.steps,
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4, "last_loss": last_loss, "losses": losses},
            "gpu_launches": launches,
            "model_tflops_per_gpu": tflops_gpu,
            "mfu_of_measured_sustained": tflops_gpu / peaks["bf16_tflops_sustained"],
            "mfu_of_measured_burst": tflops_gpu / peaks["bf16_tflops"],
        }
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def run_reference(a):
    """Hands over to baseline/run_reference.py in a fresh interpreter whose sys.path has no
    trace of this repository's packages."""
    script = os.path.join(ROOT, "baseline", "run_reference.py")
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_root, "megatron", "training.py")):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "install_reference.py")],
                           capture_output=True, text=True)
        if not os.path.isfile(os.path.join(ref_root, "megatron", "training.py")):
            if os.environ.get("RANK", "0") == "0":
                print(json.dumps({"impl": "reference", "unavailable":
                                  "reference sources not present under baseline/_ref: " + r.stderr[-200:]}))
            return 0
    tp, dp = parallel_layout(a.gpus, a.model, getattr(a, 'tp', 0), getattr(a, 'pp', 1))
    cmd = [sys.executable, script, "--model", a.model, "--gpus", str(a.gpus), "--steps", str(a.steps),
           "--warmup", str(a.warmup), "--seq", str(a.seq), "--micro_batch", str(a.micro_batch),
           "--num_microbatches", str(a.num_microbatches), "--tp", str(tp), "--dp", str(dp),
           "--debug_layers", str(a.debug_layers)]
    env = dict(os.environ)
    env.pop("PYTHONPATH", None)
    return subprocess.call(cmd, env=env, cwd=ref_root)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--model", default="llama2-7b", choices=list(MODELS))
    p.add_argument("--bucket_mb", type=float, default=256, help="A/B only: gradient bucket size")
    p.add_argument("--pp", type=int, default=1, help="other BASELINE configs only: pipeline-parallel size")
    p.add_argument("--vpp_layers", type=int, default=0, help="layers per virtual pipeline stage (interleaved 1F1B)")
    p.add_argument("--tp", type=int, default=0, help="A/B only: override the tensor-parallel size (default: 2 for N>=2)")
    p.add_argument("--seq", type=int, default=4096)
    p.add_argument("--micro_batch", type=int, default=1)
    p.add_argument("--num_microbatches", type=int, default=4)
    p.add_argument("--recompute", action="store_true")
    p.add_argument("--verbose", action="store_true")
    p.add_argument("--debug_layers", type=int, default=0, help="debug only: override the layer count")
    p.add_argument("--kernels", default="auto", help="debug only: auto|off")
    p.add_argument("--fused_comm", default="auto", help="A/B only: auto|on|off (in-kernel TP collectives)")
    p.add_argument("--fused_dp", default="auto", help="A/B only: auto|on|off (in-kernel DP reduce-scatter/all-gather)")
    p.add_argument("--fp8", action="store_true", help="A/B only (NOT the headline): FP8 hybrid linears, reported with "
                                                     "dtype fp8 and without vs_baseline")
    p.add_argument("--no_sp", action="store_true", help="A/B only: tensor parallelism without sequence parallelism")
    p.add_argument("--torch_profile", action="store_true", help="also write gpurun_out/torch_profile_nN.md")
    p.add_argument("--profile", action="store_true", help="1 warm-up + K steps, no e2e region; for ncu, NOT a bench value")
    a = p.parse_args()
    if a.warmup < 3 and not a.profile:
        a.warmup = 3
    if a.impl == "reference":
        sys.exit(run_reference(a))
    run_ours(a)


if __name__ == "__main__":
    main()
