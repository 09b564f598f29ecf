#!/bin/bash
# Logit parity of a converted checkpoint against the HF / Meta baseline.
#   bash examples/verify.sh llama2 /path/to/megatron_ckpt /path/to/baseline /path/to/data_text_document tokenizer.model
set -euo pipefail
MODEL=${1:?model}; CKPT=${2:?ckpt}; BASE=${3:?baseline}; DATA=${4:?data}; VOCAB=${5:-}
TOK=(--tokenizer_type FalconTokenizer); [[ $MODEL != falcon ]] && TOK=(--tokenizer_type SentencePieceTokenizer --vocab_file "$VOCAB")
torchrun --nproc-per-node 1 --master-addr 127.0.0.1 verify_correctness.py --model_name "$MODEL" --load "$CKPT" \
  --huggingface_cache "$BASE" --huggingface_device cuda:0 --data_path "$DATA" "${TOK[@]}" \
  --no_new_tokens --hidden_dropout 0.0 --attention_dropout 0.0 --no_bias_gelu_fusion --no_bias_dropout_fusion
