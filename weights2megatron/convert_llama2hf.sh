#!/bin/bash
# Meta Llama-1 release -> Hugging Face format (baseline for verify_correctness.py).
#   LLAMA_DIR   directory with tokenizer.model and the 7B/13B/... sub-directories of the Meta release
#   SIZE        model size in billions (7, 13, 30, 65)
#   OUT_DIR     where the Hugging Face checkpoint is written
set -euo pipefail
LLAMA_DIR=${LLAMA_DIR:?set LLAMA_DIR to the Meta release}
SIZE=${SIZE:-7}
OUT_DIR=${OUT_DIR:-$LLAMA_DIR/converted_HF_${SIZE}B}
export PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python
python3 "$(dirname "$0")/convert_llama2hf.py" --input_dir "$LLAMA_DIR" --model_size "$SIZE" --output_dir "$OUT_DIR"
