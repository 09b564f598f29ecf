#!/bin/bash
set -euo pipefail
: "${TRAIN_DATA:?}" "${VALID_DATA:?}" "${VOCAB:?}" "${PRETRAINED:?}" "${CHECKPOINT_PATH:=checkpoints/bert_mnli}"
torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tasks/main.py --task MNLI --seed 1234 \
  --train_data $TRAIN_DATA --valid_data $VALID_DATA --tokenizer_type BertWordPieceLowerCase --vocab_file "$VOCAB" \
  --epochs 5 --pretrained_checkpoint "$PRETRAINED" --num_layers 24 --hidden_size 1024 --num_attention_heads 16 \
  --micro_batch_size 8 --lr 5.0e-5 --lr_decay_style linear --lr_warmup_fraction 0.065 --seq_length 512 \
  --max_position_embeddings 512 --save_interval 500000 --save "$CHECKPOINT_PATH" --log_interval 10 \
  --eval_interval 100 --eval_iters 50 --weight_decay 1.0e-1 --bf16
