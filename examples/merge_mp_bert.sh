#!/bin/bash
# Merge a tensor-parallel BERT checkpoint into a single-rank one.
set -euo pipefail
: "${CHECKPOINT_PATH:?}" "${OUT:=${CHECKPOINT_PATH}_merged}"
python tools/checkpoint_util.py --model_type BERT --load_dir "$CHECKPOINT_PATH" --save_dir "$OUT" \
  --target_tensor_parallel_size 1 --target_pipeline_parallel_size 1
