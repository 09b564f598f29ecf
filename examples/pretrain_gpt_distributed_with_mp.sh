#!/bin/bash
# GPT pretraining, 8 GPU(s). Set DATA_PATH, VOCAB (and MERGES for GPT-2 BPE), CHECKPOINT_PATH.
set -euo pipefail
export CUDA_DEVICE_MAX_CONNECTIONS=1
: "${DATA_PATH:?}" "${VOCAB:?}" "${CHECKPOINT_PATH:=checkpoints/gpt}"
torchrun --nproc-per-node 8 --nnodes 1 --master-addr 127.0.0.1 --master-port 6000 pretrain_gpt.py \
  --num_layers 24 --hidden_size 1024 --num_attention_heads 16 --seq_length 1024 --max_position_embeddings 1024 --tensor_model_parallel_size 2 --pipeline_model_parallel_size 2 --sequence_parallel  \
  --micro_batch_size 4 --global_batch_size 64 --train_iters 500000 --lr 0.00015 --min_lr 1.0e-5 \
  --lr_decay_style cosine --lr_decay_iters 320000 --lr_warmup_fraction .01 --weight_decay 1e-2 --clip_grad 1.0 \
  --data_path $DATA_PATH --tokenizer_type GPT2BPETokenizer --vocab_file $VOCAB --merge_file $MERGES --data_impl mmap --split 949,50,1 --bf16 \
  --save $CHECKPOINT_PATH --load $CHECKPOINT_PATH --save_interval 10000 --log_interval 100 \
  --eval_interval 1000 --eval_iters 10
