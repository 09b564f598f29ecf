#!/bin/bash
# Inverse-Cloze-Task pretraining of the biencoder retriever (needs a BERT checkpoint and a titles dataset).
set -euo pipefail
: "${DATA_PATH:?}" "${TITLES_DATA_PATH:?}" "${VOCAB:?}" "${BERT_LOAD:?}" "${CHECKPOINT_PATH:=checkpoints/ict}"
torchrun --nproc-per-node 8 --master-addr 127.0.0.1 pretrain_ict.py --num_layers 12 --hidden_size 768 \
  --num_attention_heads 12 --seq_length 256 --max_position_embeddings 512 --micro_batch_size 32 \
  --train_iters 100000 --lr 0.0001 --lr_decay_style linear --lr_warmup_fraction 0.01 --weight_decay 1e-2 \
  --clip_grad 1.0 --tokenizer_type BertWordPieceLowerCase --vocab_file "$VOCAB" --data_path "$DATA_PATH" \
  --titles_data_path "$TITLES_DATA_PATH" --bert_load "$BERT_LOAD" --query_in_block_prob 0.1 \
  --biencoder_projection_dim 128 --retriever_report_topk_accuracies 1 5 20 --retriever_score_scaling \
  --save "$CHECKPOINT_PATH" --load "$CHECKPOINT_PATH" --save_interval 4000 --log_interval 100 --bf16
