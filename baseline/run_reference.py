"""Reference arm of bench.py: the UNMODIFIED reference (baseline/_ref) timed on the same
metric / config as our arm, through the reference's own public API:

    megatron.initialize.initialize_megatron -> megatron.training.setup_model_and_optimizer
    -> megatron.training.train_step(finetune.forward_step, ...)

with the reference's own model provider, batch construction, schedules, DDP, distributed
optimizer, cuBLAS GEMMs and flash-attn. What this harness supplies from OUTSIDE the
reference tree (nothing under baseline/_ref is edited):

* ``apex`` / ``amp_C`` stand-ins (baseline/shims): apex is not installed in this image;
* a synthetic token iterator and a stub tokenizer (the reference's tokenizers and datasets
  need files or network);
* two torch-2.11 compatibility shims: ``TypedStorage._untyped`` and ``torch.tensor(<storage>)``
  (used at megatron/optimizer/distrib_optimizer.py:384), and no-op fallbacks for removed
  ``torch._C._jit_*`` setters;
* ``--no_gradient_accumulation_fusion`` (apex's fused_weight_gradient_mlp_cuda is absent).

This file imports nothing from the repository's own packages.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
SHIMS = os.path.join(HERE, "shims")
REPO = os.path.dirname(HERE)

MODELS = {
    "llama2-7b": (32, 4096, 32, 32, 11008, 32000),
    "llama2-13b": (40, 5120, 40, 40, 13824, 32000),
    "llama2-70b": (80, 8192, 64, 8, 28672, 32000),
    "tiny": (2, 256, 8, 2, 688, 32000),
}


def unavailable(why):
    if os.environ.get("RANK", "0") == "0":
        print(json.dumps({"impl": "reference", "unavailable": why[:300]}), flush=True)
    sys.exit(0)


def patch_torch(torch):
    ts = torch.storage.TypedStorage
    if not hasattr(ts, "_untyped"):
        ts._untyped = lambda self: self._untyped_storage
    real_tensor = torch.tensor

    def tensor(data, *a, **kw):
        if isinstance(data, torch.UntypedStorage):
            dt = kw.get("dtype", torch.uint8)
            out = torch.empty(0, dtype=dt, device=data.device)
            out.set_(data, 0, (data.nbytes() // out.element_size(),))
            return out
        return real_tensor(data, *a, **kw)
    torch.tensor = tensor
    for name in ("_jit_set_nvfuser_enabled", "_jit_set_texpr_fuser_enabled", "_jit_set_profiling_executor",
                 "_jit_set_profiling_mode", "_jit_override_can_fuse_on_cpu", "_jit_override_can_fuse_on_gpu",
                 "_debug_set_autodiff_subgraph_inlining"):
        if not hasattr(torch._C, name):
            setattr(torch._C, name, lambda *a, **k: None)


class StubTokenizer:
    def __init__(self, vocab):
        self.vocab_size = vocab
        self.eod = vocab - 1
        self.pad = 0


class ClockSampler(threading.Thread):
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


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--model", default="llama2-7b")
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--seq", type=int, default=4096)
    p.add_argument("--micro_batch", type=int, default=1)
    p.add_argument("--num_microbatches", type=int, default=4)
    p.add_argument("--tp", type=int, default=1)
    p.add_argument("--dp", type=int, default=1)
    p.add_argument("--debug_layers", type=int, default=0)
    a = p.parse_args()

    # the reference must not see this repository's packages (it has its own `megatron`)
    sys.path[:] = [SHIMS, REF] + [x for x in sys.path if x and os.path.abspath(x) not in (REPO, HERE)]
    os.chdir(REF)
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "1")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29401")
    try:
        import torch
        patch_torch(torch)
        if not torch.cuda.is_available():
            unavailable("no CUDA device (the reference asserts CUDA, megatron/initialize.py:36)")
        import megatron  # noqa: F401
        assert os.path.abspath(megatron.__file__).startswith(REF), megatron.__file__
        import finetune as ref_finetune
        import megatron.global_vars as gv
        import megatron.training as ref_training
        from megatron import get_args
        from megatron.initialize import initialize_megatron
        from megatron.model import ModelType
        from megatron.core import mpu
    except SystemExit:
        raise
    except Exception as e:  # noqa: BLE001
        unavailable(f"reference import failed: {type(e).__name__}: {e}")

    L, h, nh, nkv, ffn, vocab = MODELS[a.model]
    if a.debug_layers:
        L = a.debug_layers
    gbs = a.micro_batch * a.num_microbatches * a.dp
    argv = [
        "--model_name", "llama2", "--num_layers", str(L), "--hidden_size", str(h),
        "--num_attention_heads", str(nh), "--num_attention_heads_kv", str(nkv),
        "--ffn_hidden_size", str(ffn), "--seq_length", str(a.seq), "--max_position_embeddings", str(a.seq),
        "--micro_batch_size", str(a.micro_batch), "--global_batch_size", str(gbs),
        "--train_iters", str(a.steps * 2 + a.warmup + 8), "--lr", "1e-5", "--min_lr", "1e-6",
        "--lr_decay_style", "cosine", "--lr_warmup_iters", "1", "--weight_decay", "0.1",
        "--adam_beta2", "0.95", "--clip_grad", "1.0",
        "--use_rms_norm", "--glu_activation", "swiglu", "--no_tie_embed_logits",
        "--position_embedding_type", "rotary", "--use_flash_attn", "--no_bias_gelu_fusion",
        "--hidden_dropout", "0.0", "--attention_dropout", "0.0", "--bf16",
        "--tensor_model_parallel_size", str(a.tp), "--pipeline_model_parallel_size", "1",
        "--use_distributed_optimizer", "--no_gradient_accumulation_fusion",
        "--tokenizer_type", "SentencePieceTokenizer", "--vocab_file", "synthetic",
        "--make_vocab_size_divisible_by", "128",
        "--num_workers", "0", "--log_interval", "1000000", "--eval_iters", "0",
        "--eval_interval", "1000000",
    ]
    if a.tp > 1:
        argv.append("--sequence_parallel")
    sys.argv = ["finetune.py"] + argv

    # synthetic-data hook: stub tokenizer, padded vocab set by hand
    def fake_build_tokenizer(args):
        args.padded_vocab_size = vocab
        gv._GLOBAL_TOKENIZER = StubTokenizer(vocab)
        return gv._GLOBAL_TOKENIZER
    gv._build_tokenizer = fake_build_tokenizer

    rank0 = os.environ.get("RANK", "0") == "0"
    devnull = open(os.devnull, "w")
    real_stdout = sys.stdout
    sys.stdout = devnull
    try:
        initialize_megatron(ref_finetune.extra_args, {})
        args = get_args()
        model, optimizer, sched = ref_training._setup_model_and_optimizer(
            ref_finetune.model_provider, ModelType.encoder_or_decoder, args=args)
    except Exception as e:  # noqa: BLE001
        sys.stdout = real_stdout
        import traceback
        traceback.print_exc()
        unavailable(f"reference set-up failed: {type(e).__name__}: {e}")
    for m in model:
        m.train()
    import torch.distributed as dist
    dev = torch.device("cuda", torch.cuda.current_device())
    gen = torch.Generator().manual_seed(1234 + mpu.get_data_parallel_rank())
    pool = [torch.randint(0, vocab, (a.micro_batch, a.seq + 1), generator=gen, dtype=torch.int64).pin_memory()
            for _ in range(8)]

    def batches():
        i = 0
        while True:
            yield {"text": pool[i % len(pool)]}
            i += 1
    it = batches() if mpu.get_tensor_model_parallel_rank() == 0 else None
    n_mb = gbs // (a.micro_batch * a.dp)
    h2d = n_mb * pool[0].numel() * pool[0].element_size()

    def step():
        return ref_training.train_step(ref_finetune.forward_step, it, model, optimizer, sched, args)

    try:
        dbg = []
        for _ in range(max(a.warmup, 3)):
            r = step()
            if os.environ.get('B200_BENCH_DEBUG_LOSSES') and r[0]:
                dbg.append(float(r[0]['lm loss'].item()))
        if dbg and rank0:
            print('DEBUG warmup losses', dbg, file=sys.stderr, flush=True)
        torch.cuda.synchronize()
        dist.barrier()
        sampler = ClockSampler(torch.cuda.current_device())
        sampler.start()
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
        t0 = time.perf_counter()
        last, losses = None, []
        for _ in range(a.steps):
            loss_dict = step()[0]
            if loss_dict:
                last = float(loss_dict["lm loss"].item())
                losses.append(last)
            else:
                torch.cuda.current_stream().synchronize()
        torch.cuda.synchronize()
        dist.barrier()
        e2e_s = time.perf_counter() - t0
        clocks = sampler.stop()
    except Exception as e:  # noqa: BLE001
        sys.stdout = real_stdout
        import traceback
        traceback.print_exc()
        unavailable(f"reference step failed: {type(e).__name__}: {e}")
    t = torch.tensor([ms, e2e_s * 1000.0], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sys.stdout = real_stdout
    tokens = gbs * a.seq
    if rank0:
        print(json.dumps({
            "impl": "reference",
            "metric": "Llama-2 7B bf16 training throughput (tokens/s, whole job, max over ranks)",
            "value": tokens * a.steps / (float(t[0]) / 1000.0), "unit": "tokens/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": float(t[0]) / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic tokens, random-init weights",
            "config": {"model": a.model, "global_batch": gbs, "seq_len": a.seq, "micro_batch": a.micro_batch,
                       "num_microbatches": n_mb, "parallelism": f"tp{a.tp}" + (f"dp{a.dp}" if a.dp > 1 else ""),
                       "sequence_parallel": a.tp > 1, "distributed_optimizer": True,
                       "substitutions": "apex FusedAdam -> torch AdamW(fused); amp_C -> torch._foreach; "
                                        "no wgrad-accumulation fusion; synthetic tokens; torch-2.11 storage shim"},
            "clocks": clocks,
            "e2e": {"value": tokens * a.steps / (float(t[1]) / 1000.0), "unit": "tokens/s",
                    "ms_per_step": float(t[1]) / a.steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "last_loss": last, "losses": losses},
        }), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
