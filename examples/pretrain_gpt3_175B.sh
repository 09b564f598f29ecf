#!/bin/bash
# GPT-3 175B layout (SLURM, 128 nodes x 8 GPUs): tp 8, pp 16, interleaved schedule, batch ramp-up.
#SBATCH --nodes=128 --ntasks-per-node=8 --exclusive
set -euo pipefail
export CUDA_DEVICE_MAX_CONNECTIONS=1
: "${DATA_PATH:?}" "${VOCAB:?}" "${MERGES:?}" "${CHECKPOINT_PATH:=checkpoints/gpt3_175b}"
srun python pretrain_gpt.py --tensor_model_parallel_size 8 --pipeline_model_parallel_size 16 \
  --num_layers_per_virtual_pipeline_stage 1 --sequence_parallel --use_distributed_optimizer \
  --num_layers 96 --hidden_size 12288 --num_attention_heads 96 --seq_length 2048 --max_position_embeddings 2048 \
  --micro_batch_size 1 --global_batch_size 1536 --rampup_batch_size 16 16 5859375 --train_samples 146484375 \
  --lr_decay_samples 126953125 --lr_warmup_samples 183105 --lr 6.0e-5 --min_lr 6.0e-6 --lr_decay_style cosine \
  --clip_grad 1.0 --weight_decay 0.1 --adam_beta1 0.9 --adam_beta2 0.95 --init_method_std 0.006 \
  --tokenizer_type GPT2BPETokenizer --vocab_file "$VOCAB" --merge_file "$MERGES" --data_path "$DATA_PATH" \
  --data_impl mmap --split 949,50,1 --bf16 --recompute_granularity full --recompute_method uniform \
  --save "$CHECKPOINT_PATH" --load "$CHECKPOINT_PATH" --save_interval 1000 --log_interval 10 \
  --eval_interval 1000 --eval_iters 40
