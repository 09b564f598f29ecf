#!/bin/bash
# Paths and common options of the scaling sweep. Edit for your cluster.
export SC21_ROOT=${SC21_ROOT:-$PWD}
export DATA_PATH=${DATA_PATH:-$SC21_ROOT/data/gpt2_text_document}
export VOCAB=${VOCAB:-$SC21_ROOT/data/gpt2-vocab.json}
export MERGES=${MERGES:-$SC21_ROOT/data/gpt2-merges.txt}
export LOG_DIR=${LOG_DIR:-$SC21_ROOT/logs}
export COMMON="--seq_length 2048 --max_position_embeddings 2048 --train_iters 50 --lr 1.0e-4 --min_lr 1.0e-5 \
 --lr_decay_style cosine --tokenizer_type GPT2BPETokenizer --vocab_file $VOCAB --merge_file $MERGES \
 --data_path $DATA_PATH --split 949,50,1 --bf16 --log_interval 1 --eval_iters 0 --log_throughput \
 --recompute_granularity full --recompute_method uniform"
