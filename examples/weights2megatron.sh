#!/bin/bash
# HF / Meta weights -> single-rank Megatron "release" checkpoint.
#   bash examples/weights2megatron.sh llama2 7 /path/to/weights /path/to/out [/path/to/tokenizer.model]
set -euo pipefail
MODEL=${1:?model}; SIZE=${2:?size}; SRC=${3:?source dir}; OUT=${4:?out dir}; VOCAB=${5:-}
ARGS=(); [[ -n $VOCAB ]] && ARGS=(--vocab_file "$VOCAB")
python weights2megatron/weights2megatron.py "$MODEL" --size "$SIZE" --out "$OUT" --cache_dir "$SRC" "${ARGS[@]}"
