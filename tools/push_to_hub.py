"""Load a Hugging Face checkpoint written by weights2megatron/megatron2hf.py, optionally recast
it / stamp RoPE scaling, and push it to the Hub (tools/push_to_hub.py:8-161 surface).
Needs network access and a token; everything before the push also works offline with
``--dry_run``."""
import argparse
import os


def parse():
    p = argparse.ArgumentParser(description="Push a converted model to the Hugging Face Hub")
    p.add_argument("model_name", help="path of the HF checkpoint directory")
    p.add_argument("--hf_repo_name", type=str, required=False, help="<org>/<repo> to push to")
    p.add_argument("--dtype", default="fp16", choices=["fp16", "bf16", "fp32"])
    p.add_argument("--auth_token", type=str, default=os.environ.get("HF_TOKEN"))
    p.add_argument("--output_folder", type=str, default=None, help="also save the recast model here")
    p.add_argument("--max_shard_size", type=str, default="10GB")
    p.add_argument("--unsafe", action="store_true", help="save .bin instead of safetensors")
    p.add_argument("--rope_scaling_type", type=str, default=None, choices=[None, "linear", "dynamic"])
    p.add_argument("--rope_scaling_factor", type=float, default=None)
    p.add_argument("--trust_remote_code", action="store_true")
    p.add_argument("--dry_run", action="store_true", help="do everything except the upload")
    return p.parse_args()


def main():
    import torch
    import transformers
    a = parse()
    dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[a.dtype]
    cfg = transformers.AutoConfig.from_pretrained(a.model_name, trust_remote_code=a.trust_remote_code)
    if a.rope_scaling_type is not None:
        assert a.rope_scaling_factor and a.rope_scaling_factor >= 1.0, "--rope_scaling_factor >= 1 required"
        cfg.rope_scaling = {"type": a.rope_scaling_type, "factor": a.rope_scaling_factor}
        print("rope_scaling ->", cfg.rope_scaling)
    tok = transformers.AutoTokenizer.from_pretrained(a.model_name, trust_remote_code=a.trust_remote_code)
    model = transformers.AutoModelForCausalLM.from_pretrained(
        a.model_name, config=cfg, torch_dtype=dtype, trust_remote_code=a.trust_remote_code)
    print(f"loaded {sum(p.numel() for p in model.parameters()) / 1e9:.2f} B parameters as {dtype}")
    if a.output_folder:
        model.save_pretrained(a.output_folder, max_shard_size=a.max_shard_size,
                              safe_serialization=not a.unsafe)
        tok.save_pretrained(a.output_folder)
    if a.dry_run or not a.hf_repo_name:
        print("dry run: nothing pushed")
        return
    model.push_to_hub(a.hf_repo_name, token=a.auth_token, max_shard_size=a.max_shard_size,
                      safe_serialization=not a.unsafe)
    tok.push_to_hub(a.hf_repo_name, token=a.auth_token)


if __name__ == "__main__":
    main()
