"""Knowledge / response generation by few-shot prompting (tasks/msdp/prompt.py:19-308 surface),
either against a served model (REST) or a locally loaded checkpoint."""
import json

import torch

from epfl_megatron_b200 import get_args, get_tokenizer, print_rank_0
from epfl_megatron_b200.checkpointing import load_checkpoint
from epfl_megatron_b200.models import FalconModel, GPTModel, LlamaModel
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.text_generation import generate_and_post_process
from epfl_megatron_b200.training import get_model


def call_model_api(inputs, tokens_to_generate):
    """One prompt -> the generated continuation, through PUT <url>/api of a running server."""
    import urllib.request
    args = get_args()
    req = urllib.request.Request(args.megatron_api_url, method="PUT",
                                 data=json.dumps({"prompts": [inputs], "tokens_to_generate": tokens_to_generate,
                                                  "top_k": 1}).encode(),
                                 headers={"Content-Type": "application/json; charset=UTF-8"})
    with urllib.request.urlopen(req) as r:
        text = json.loads(r.read().decode())["text"][0]
    return text[len(inputs):].split("\n")[0].strip()


def read_prompts(prompt_path, prompt_type, n_example):
    """knowledge: {topic + last turn -> [examples]} per test sample; response: one fixed prompt."""
    if prompt_type == "knowledge":
        prompts = {}
        with open(prompt_path, encoding="utf-8") as f:
            for line in f:
                d = json.loads(line.strip())
                for key, examples in d.items():
                    prompts[key] = "".join(e.strip() + " \n" for e in examples[:n_example])
        return prompts
    with open(prompt_path, encoding="utf-8") as f:
        lines = [ln.strip() for ln in f.readlines()[:n_example]]
    return "".join(ln + " \n" for ln in lines)


def build_input(line, prompts, prompt_type):
    splits = line.strip().split("\t")
    topic = splits[0]
    if prompt_type == "knowledge":
        turns = splits[1].split(" [SEP] ")
        last = turns[-1]
        key = topic + " " + last
        return prompts[key] + "( " + last + " ) " + topic + " =>"
    turns = splits[1].split(" [SEP] ")
    knowledge = splits[2]
    last = " ".join(turns[-1].split())
    knowledge = " ".join(knowledge.split())
    return prompts + "Topic: " + topic + ". User says: " + last + " We know that: " + knowledge + \
        " System replies:"


def generate_samples_by_calling_api():
    args = get_args()
    assert args.prompt_type in ("knowledge", "response")
    prompts = read_prompts(args.prompt_file, args.prompt_type, args.num_prompt_examples)
    with open(args.sample_input_file, encoding="utf-8") as fin, \
            open(args.sample_output_file, "w", encoding="utf-8") as fout:
        for n, line in enumerate(fin, start=1):
            out = call_model_api(build_input(line, prompts, args.prompt_type), args.out_seq_length)
            fout.write(out + "\n")
            if n % 10 == 0:
                print(f"processed {n} samples", flush=True)


def model_provider(pre_process=True, post_process=True):
    args = get_args()
    print_rank_0("building GPT model ...")
    cls = {"falcon": FalconModel, "llama": LlamaModel, "llama2": LlamaModel, "codellama": LlamaModel}.get(
        getattr(args, "model_name", "gpt"), GPTModel)
    return cls(num_tokentypes=0, parallel_output=False, pre_process=pre_process, post_process=post_process)


def generate_samples_by_prompting_input_from_file(model):
    args = get_args()
    assert args.sample_input_file is not None, "sample input file is not provided."
    first = mpu.is_pipeline_first_stage() and mpu.get_tensor_model_parallel_rank() == 0
    lines, fout, prompts = [], None, None
    if first:
        with open(args.sample_input_file, encoding="utf-8") as f:
            lines = f.readlines()
        out_path = args.sample_output_file or args.sample_input_file + ".out"
        fout = open(out_path, "w", encoding="utf-8")
        prompts = read_prompts(args.prompt_file, args.prompt_type, args.num_prompt_examples)
    n = torch.tensor([len(lines)], dtype=torch.int64)
    if torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        n = n.to(dev)
        torch.distributed.broadcast(n, 0)
    model.eval()
    with torch.no_grad():
        for i in range(int(n.item())):
            raw = build_input(lines[i], prompts, args.prompt_type) if first else None
            out = generate_and_post_process(model, prompts=[raw] if first else None,
                                            tokens_to_generate=args.out_seq_length, top_k_sampling=1)
            if first:
                text = out[0][0][len(raw):].split("\n")[0].strip()
                fout.write(text + "\n")
                if (i + 1) % 10 == 0:
                    print(f"processed {i + 1} samples", flush=True)
    if fout is not None:
        fout.close()


def main():
    args = get_args()
    if args.api_prompt:
        generate_samples_by_calling_api()
        return
    if args.num_layers_per_virtual_pipeline_stage is not None:
        print("Interleaved pipeline schedule is not yet supported for text generation.")
        return
    model = get_model(model_provider, wrap_with_ddp=False)
    if args.load is not None:
        load_checkpoint(model, None, None)
    assert len(model) == 1
    generate_samples_by_prompting_input_from_file(model[0])
