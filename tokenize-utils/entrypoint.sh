#!/bin/bash
# Runs inside the container (see run_dock.sh): tokenise the train / validation jsonl files found
# under $DATA_DIR into indexed datasets next to them.
set -euo pipefail
REPO=${REPO:-/opt/megatron}
DATA_DIR=${DATA_DIR:-/data}
TOKENIZER=${TOKENIZER:-SentencePieceTokenizer}
VOCAB=${VOCAB:-$DATA_DIR/tokenizer.model}
cd "$REPO"
python -m epfl_megatron_b200.ops.build --data-only
for split in train valid; do
  [ -f "$DATA_DIR/$split.json" ] || continue
  python tools/preprocess_data.py --input "$DATA_DIR/$split.json" --output_prefix "$DATA_DIR/$split" \
    --dataset_impl mmap --tokenizer_type "$TOKENIZER" --vocab_file "$VOCAB" --workers "${WORKERS:-2}" \
    --chunk_size 2048 --append_eod
done
