#!/bin/bash
# Serve a 345M GPT checkpoint with 8-way tensor parallelism: PUT http://host:5000/api  {"prompts": [...], "tokens_to_generate": N}
set -euo pipefail
: "${CHECKPOINT:?}" "${VOCAB:?}" "${MERGES:?}"
export CUDA_DEVICE_MAX_CONNECTIONS=1
torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tools/run_text_generation_server.py \
  --tensor_model_parallel_size 8 --pipeline_model_parallel_size 1 --num_layers 24 --hidden_size 1024 \
  --load "$CHECKPOINT" --num_attention_heads 16 --max_position_embeddings 1024 --tokenizer_type GPT2BPETokenizer \
  --bf16 --micro_batch_size 1 --seq_length 1024 --vocab_file "$VOCAB" --merge_file "$MERGES" --seed 42
