#!/bin/bash
# Finetune / continue pretraining of Llama, Llama-2, CodeLlama or Falcon on one 8 x B200 node.
#   bash examples/finetune.sh llama2 --size 7 --tp 2 --pp 1 --micro-batch 1 --global-batch 64 \
#        --data /path/to/tokenized_text_document --vocab /path/to/tokenizer.model --checkpoint /path/to/ckpt
set -euo pipefail
MODEL=${1:?usage: finetune.sh <llama|llama2|codellama|falcon|gpt> [options]}; shift
SIZE=7; TP=2; PP=1; GPUS=8; MICRO=1; GLOBAL=64; ITERS=1000; SEQ=""; LR=3e-4
DATA=""; VOCAB=""; CKPT=""; OUT=checkpoints/run; RANK=0; NODES=1; ADDR=127.0.0.1; PORT=6000; EXTRA=()
while [[ $# -gt 0 ]]; do case $1 in
  --size) SIZE=$2; shift 2;; --tp) TP=$2; shift 2;; --pp) PP=$2; shift 2;; --gpus) GPUS=$2; shift 2;;
  --micro-batch) MICRO=$2; shift 2;; --global-batch) GLOBAL=$2; shift 2;; --iters) ITERS=$2; shift 2;;
  --seq-len) SEQ=$2; shift 2;; --lr) LR=$2; shift 2;; --data) DATA=$2; shift 2;; --vocab) VOCAB=$2; shift 2;;
  --checkpoint) CKPT=$2; shift 2;; --out) OUT=$2; shift 2;; --rank) RANK=$2; shift 2;; --nodes) NODES=$2; shift 2;;
  --addr) ADDR=$2; shift 2;; --port) PORT=$2; shift 2;; *) EXTRA+=("$1"); shift;; esac; done
if [[ $MODEL == falcon ]]; then
  TOK=(--tokenizer_type FalconTokenizer); SEQ=${SEQ:-2048}
  ARCH=(--parallel_attn --position_embedding_type rotary --use_flash_attn --no_bias_gelu_fusion)
  [[ $SIZE == 40 ]] && ARCH+=(--parallel_layernorm)
else
  TOK=(--tokenizer_type SentencePieceTokenizer --vocab_file "$VOCAB"); SEQ=${SEQ:-4096}
  [[ $MODEL == llama ]] && SEQ=${SEQ:-2048}
  ARCH=(--use_rms_norm --glu_activation swiglu --no_tie_embed_logits --position_embedding_type rotary
        --use_flash_attn --no_bias_gelu_fusion --layernorm_epsilon 1e-5 --hidden_dropout 0.0 --attention_dropout 0.0)
fi
LOAD=(); [[ -n $CKPT ]] && LOAD=(--load "$CKPT" --use_checkpoint_args --finetune)
DATA_ARGS=(--synthetic_data); [[ -n $DATA ]] && DATA_ARGS=(--data_path "$DATA" --split 989,10,1)
SP=(); [[ $TP -gt 1 ]] && SP=(--sequence_parallel)
export CUDA_DEVICE_MAX_CONNECTIONS=1
torchrun --nproc-per-node "$GPUS" --nnodes "$NODES" --node-rank "$RANK" --master-addr "$ADDR" --master-port "$PORT" \
  finetune.py --model_name "$MODEL" "${ARCH[@]}" "${TOK[@]}" "${LOAD[@]}" "${DATA_ARGS[@]}" "${SP[@]}" \
  --tensor_model_parallel_size "$TP" --pipeline_model_parallel_size "$PP" --use_distributed_optimizer --bf16 \
  --seq_length "$SEQ" --max_position_embeddings "$SEQ" --micro_batch_size "$MICRO" --global_batch_size "$GLOBAL" \
  --train_iters "$ITERS" --lr "$LR" --min_lr 3e-5 --lr_decay_style cosine --lr_warmup_iters 100 \
  --weight_decay 0.1 --adam_beta1 0.9 --adam_beta2 0.95 --clip_grad 1.0 --recompute_granularity selective \
  --save "$OUT" --save_interval 500 --log_interval 10 --eval_interval 500 --eval_iters 10 \
  --log_throughput --metrics_jsonl "$OUT/metrics.jsonl" "${EXTRA[@]}"
