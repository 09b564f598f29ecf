#!/bin/bash
# The five BASELINE.json configurations through bench.py (synthetic tokens, random-init weights) on one
# 8 x B200 node. Each prints one JSON line (tokens/s whole job, device-timed, max over ranks).
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"

# 2: Llama-2 7B, TP2 x DP4, sequence parallel, distributed optimizer (the headline config)
$RUN --master-port 29501 bench.py --gpus 8 --steps 10 --warmup 3
# 3: Llama-2 13B, TP4 x PP2, interleaved 1F1B with 4 chunks of 5 layers per stage
$RUN --master-port 29502 bench.py --gpus 8 --model llama2-13b --tp 4 --pp 2 --vpp_layers 5 --steps 5 --warmup 3
# 4: Falcon-40B, TP8, parallel attention / MLP, sequence 2048
$RUN --master-port 29503 bench.py --gpus 8 --model falcon-40b --tp 8 --seq 2048 --steps 5 --warmup 3
# 5: Llama-2 70B, TP4 x PP2, full activation recompute, distributed optimizer (~155 GB of state per GPU)
PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True \
$RUN --master-port 29504 bench.py --gpus 8 --model llama2-70b --tp 4 --pp 2 --recompute --steps 3 --warmup 3
# 1: tiny Llama on CPU / Gloo, TP2 (plumbing only): python -m pytest tests/test_pipeline_cpu.py -k tp2
