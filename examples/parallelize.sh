#!/bin/bash
# Re-shard a checkpoint: bash examples/parallelize.sh <model> <in_dir> <out_dir> <tp> <pp> [true_vocab_size]
set -euo pipefail
MODEL=${1:?model}; IN=${2:?in}; OUT=${3:?out}; TP=${4:-1}; PP=${5:-1}; V=${6:-}
ARGS=(); [[ -n $V ]] && ARGS=(--true_vocab_size "$V")
python tools/checkpoint_util.py --model_type "$MODEL" --load_dir "$IN" --save_dir "$OUT" \
  --target_tensor_parallel_size "$TP" --target_pipeline_parallel_size "$PP" --bf16 "${ARGS[@]}"
