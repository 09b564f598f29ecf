"""Runs a few training steps of a small Llama through the public API and prints the
per-step losses as JSON. Used by the parity tests (CUDA kernels vs PyTorch restatement,
TP=1 vs TP=2, ...) which launch it in subprocesses / under torchrun."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--kernels", default="auto")
    p.add_argument("--gemm", default="auto")
    p.add_argument("--tp", type=int, default=1)
    p.add_argument("--pp", type=int, default=1)
    p.add_argument("--vpp_layers", type=int, default=0)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--dtype", default="bf16")
    p.add_argument("--hidden", type=int, default=1024)
    p.add_argument("--heads", type=int, default=8)
    p.add_argument("--kv_heads", type=int, default=2)
    p.add_argument("--layers", type=int, default=2)
    p.add_argument("--seq", type=int, default=256)
    p.add_argument("--model_name", default="llama2")
    p.add_argument("--dist_opt", action="store_true")
    p.add_argument("--recompute", action="store_true")
    p.add_argument("--backend", default=None)
    p.add_argument("--lr", type=float, default=1e-3)
    p.add_argument("--save", default=None)
    p.add_argument("--load", default=None)
    p.add_argument("--num_microbatches", type=int, default=2)
    p.add_argument("--no_overlap", action="store_true")
    p.add_argument("--bucket_mb", type=int, default=None)
    p.add_argument("--fused_comm", default="auto")
    p.add_argument("--fused_dp", default="auto")
    p.add_argument("--check_dp", type=int, default=0)
    p.add_argument("--corrupt_rank", type=int, default=-1, help="test hook: perturb one parameter on this rank")
    p.add_argument("--nvtx", action="store_true")
    p.add_argument("--no_sp", action="store_true", help="tensor parallel without sequence parallelism")
    p.add_argument("--extra", default="", help="more training flags, space separated")
    a = p.parse_args()

    import torch
    import finetune
    from epfl_megatron_b200 import get_args
    from epfl_megatron_b200.initialize import initialize_megatron
    from epfl_megatron_b200.models.enums import ModelType
    from epfl_megatron_b200.parallel import state as mpu
    from epfl_megatron_b200.training import setup_model_and_optimizer, train_step
    from epfl_megatron_b200.checkpointing import save_checkpoint

    world = int(os.environ.get("WORLD_SIZE", "1"))
    dp = world // (a.tp * a.pp)
    mbs = 2
    argv = ["--model_name", a.model_name, "--num_layers", str(a.layers), "--hidden_size", str(a.hidden),
            "--num_attention_heads", str(a.heads), "--num_attention_heads_kv", str(a.kv_heads),
            "--seq_length", str(a.seq), "--max_position_embeddings", str(a.seq),
            "--micro_batch_size", str(mbs), "--global_batch_size", str(mbs * a.num_microbatches * dp),
            "--train_iters", "100", "--lr", str(a.lr), "--min_lr", str(a.lr / 10), "--lr_decay_style", "constant",
            "--position_embedding_type", "rotary", "--use_flash_attn", "--no_bias_gelu_fusion",
            "--hidden_dropout", "0.0", "--attention_dropout", "0.0",
            "--tensor_model_parallel_size", str(a.tp), "--pipeline_model_parallel_size", str(a.pp),
            "--tokenizer_type", "NullTokenizer", "--synthetic_data", "--num_workers", "0",
            "--log_interval", "1000", "--eval_iters", "0", "--b200_kernels", a.kernels, "--b200_gemm", a.gemm,
            "--use_cpu_initialization", "--make_vocab_size_divisible_by", "64", "--clip_grad", "1.0"]
    if a.model_name.startswith("llama"):
        argv += ["--use_rms_norm", "--glu_activation", "swiglu", "--no_tie_embed_logits",
                 "--ffn_hidden_size", str(int(a.hidden * 2.75) // 64 * 64)]
    elif a.model_name == "falcon":
        argv += ["--parallel_attn", "--parallel_layernorm"]
    if a.dtype == "bf16":
        argv.append("--bf16")
    if a.tp > 1 and not a.no_sp:
        argv.append("--sequence_parallel")
    argv += a.extra.split()
    if a.dist_opt:
        argv.append("--use_distributed_optimizer")
    if a.no_overlap:
        argv.append("--no_overlap_grad_reduce")
    if a.bucket_mb is not None:
        argv += ["--ddp_bucket_mb", str(a.bucket_mb)]
    argv += ["--b200_fused_comm", a.fused_comm, "--b200_fused_dp", a.fused_dp]
    if a.check_dp:
        argv += ["--check_dp_consistency_interval", str(a.check_dp)]
    if a.nvtx:
        argv.append("--nvtx_ranges")
    if a.recompute:
        argv += ["--recompute_granularity", "full", "--recompute_method", "uniform"]
    if a.vpp_layers:
        argv += ["--num_layers_per_virtual_pipeline_stage", str(a.vpp_layers)]
    if a.backend:
        argv += ["--distributed_backend", a.backend]
    if a.save:
        argv += ["--save", a.save, "--save_interval", "1000"]
    if a.load:
        argv += ["--load", a.load]
    real = sys.stdout
    sys.stdout = open(os.devnull, "w")
    initialize_megatron(finetune.extra_args, {}, args_list=argv)
    args = get_args()
    model, optimizer, sched = setup_model_and_optimizer(finetune.model_provider, ModelType.encoder_or_decoder)
    for m in model:
        m.train()
    gen = torch.Generator().manual_seed(77 + mpu.get_data_parallel_rank())
    pool = [torch.randint(0, 512, (mbs, a.seq + 1), generator=gen) for _ in range(4)]

    def it():
        i = 0
        while True:
            yield {"text": pool[i % 4]}
            i += 1
    data = it()
    if args.virtual_pipeline_model_parallel_size is not None:
        data = [it() for _ in range(len(model))]
    losses, norms = [], []
    for s in range(a.steps):
        args.curr_iteration = s
        if s == 1 and a.corrupt_rank == int(os.environ.get("RANK", "0")):
            with torch.no_grad():
                next(model[0].parameters()).view(-1)[3] += 1.0
        ld, skipped, gn, _ = train_step(finetune.forward_step, data, model, optimizer, sched, args)
        if ld:
            losses.append(float(ld["lm loss"].item()))
        norms.append(float(gn) if gn is not None else None)
    if a.save:
        save_checkpoint(args.iteration + a.steps, model, optimizer, sched)
    fused_dp_used = False
    for ch in getattr(optimizer, "chunks", []):
        if getattr(ch, "ws", None) is not None:
            ch.ws.check()          # raises if an in-kernel cross-GPU wait timed out
            fused_dp_used = True
    sys.stdout = real
    import torch.distributed as dist
    if mpu.is_pipeline_last_stage(ignore_virtual=True) and mpu.get_tensor_model_parallel_rank() == 0 \
            and mpu.get_data_parallel_rank() == 0:
        print("RESULT " + json.dumps({"losses": losses, "grad_norms": norms, "fused_dp": fused_dp_used}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
