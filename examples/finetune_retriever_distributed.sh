#!/bin/bash
set -euo pipefail
: "${TRAIN_DATA:?}" "${VALID_DATA:?}" "${VOCAB:?}" "${PRETRAINED:?}" "${EVIDENCE:?}" "${CHECKPOINT_PATH:=checkpoints/retriever}"
torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tasks/main.py --task RET-FINETUNE-NQ --train_with_neg \
  --train_hard_neg 1 --pretrained_checkpoint "$PRETRAINED" --num_layers 12 --hidden_size 768 \
  --num_attention_heads 12 --tokenizer_type BertWordPieceLowerCase --vocab_file "$VOCAB" --epochs 80 \
  --micro_batch_size 8 --eval_micro_batch_size 16 --indexer_batch_size 128 --train_data $TRAIN_DATA \
  --valid_data $VALID_DATA --evidence_data_path "$EVIDENCE" --seq_length 512 --retriever_seq_length 256 \
  --max_position_embeddings 512 --retriever_score_scaling --retriever_report_topk_accuracies 1 5 10 20 100 \
  --lr 2e-5 --lr_decay_style linear --lr_warmup_fraction 0.01 --weight_decay 1e-1 --clip_grad 2.0 \
  --save "$CHECKPOINT_PATH" --save_interval 5000 --log_interval 10 --eval_interval 25000 --eval_iters 100 --bf16
