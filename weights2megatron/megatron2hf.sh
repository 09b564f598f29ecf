#!/bin/bash
# Megatron checkpoint (any TP x PP sharding) -> Hugging Face model + tokenizer.
#   INPUT_DIR    Megatron checkpoint directory (with latest_checkpointed_iteration.txt)
#   OUTPUT_DIR   where the Hugging Face checkpoint is written
#   MODEL        llama | llama2 | codellama | falcon
#   SHARDED      1 if INPUT_DIR holds more than one mp_rank_* shard (it is merged to TP1 x PP1 first)
#   VOCAB_FILE   tokenizer.model of the run (Llama family); TRUE_VOCAB_SIZE strips the vocabulary padding
set -euo pipefail
INPUT_DIR=${INPUT_DIR:?set INPUT_DIR to the Megatron checkpoint}
OUTPUT_DIR=${OUTPUT_DIR:?set OUTPUT_DIR}
MODEL=${MODEL:-llama2}
SHARDED=${SHARDED:-0}
TRUE_VOCAB_SIZE=${TRUE_VOCAB_SIZE:-32000}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
export PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python
cd "$ROOT"
SRC=$INPUT_DIR
if [ "$SHARDED" -gt 0 ]; then
    SRC=$(mktemp -d /tmp/unsharded.XXXXXX)
    python3 tools/checkpoint_util.py --target_tensor_parallel_size 1 --target_pipeline_parallel_size 1 \
        --load_dir "$INPUT_DIR" --save_dir "$SRC" --model_type "$MODEL" --true_vocab_size "$TRUE_VOCAB_SIZE" --bf16
fi
python3 weights2megatron/megatron2hf.py --input_dir "$SRC" --output_dir "$OUTPUT_DIR" --model "$MODEL" \
    ${VOCAB_FILE:+--vocab_file "$VOCAB_FILE"} --true_vocab_size "$TRUE_VOCAB_SIZE"
if [ "$SHARDED" -gt 0 ]; then rm -r "$SRC"; fi
