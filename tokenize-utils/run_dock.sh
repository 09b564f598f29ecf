#!/bin/bash
# Build the tokenisation image and run entrypoint.sh in it with the data directory mounted.
set -euo pipefail
DATA_DIR=${1:?host directory with train.json / valid.json and the tokenizer model}
docker build -t megatron-b200-tokenize -f "$(dirname "$0")/Dockerfile" "$(dirname "$0")/.."
docker run --rm --shm-size=2gb -v "$DATA_DIR":/data --entrypoint bash megatron-b200-tokenize tokenize-utils/entrypoint.sh
