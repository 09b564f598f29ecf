#!/bin/bash
set -euo pipefail
: "${TASK:=LAMBADA}" "${VALID_DATA:?}" "${VOCAB:?}" "${MERGES:?}" "${CHECKPOINT_PATH:?}"
torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tasks/main.py --task "$TASK" --valid_data "$VALID_DATA" \
  --tokenizer_type GPT2BPETokenizer --strict_lambada --vocab_file "$VOCAB" --merge_file "$MERGES" \
  --load "$CHECKPOINT_PATH" --tensor_model_parallel_size 1 --num_layers 24 --hidden_size 1024 \
  --num_attention_heads 16 --micro_batch_size 8 --seq_length 1024 --max_position_embeddings 1024 \
  --log_interval 10 --bf16 --no_load_optim --no_load_rng
