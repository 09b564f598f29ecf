#!/bin/bash
set -euo pipefail
IN=${1:?input jsonl}; OUT=${2:?output prefix}; VOCAB=${3:?tokenizer.model}; WORKERS=${WORKERS:-16}
python "$(dirname "$0")/../tools/preprocess_data.py" --input "$IN" --output_prefix "$OUT" \
  --tokenizer_type SentencePieceTokenizer --vocab_file "$VOCAB" --workers "$WORKERS" --chunk_size 32 --append_eod
