#!/bin/bash
# Micro batch size sweep: bash run_figure_16.sh <MBS: 1|2|4|8> <GBS: 128|512>   (TP 8 x PP 8, h 15360)
set -euo pipefail
. "$(dirname "$0")/CONFIG.sh"
MBS=${1:?micro batch size}; GBS=${2:-128}
TP=8; PP=8; NLS=32; HS=15360; NAH=128; NNODES=8
export NNODES
export OPTIONS="--tensor_model_parallel_size $TP --pipeline_model_parallel_size $PP --micro_batch_size $MBS \
 --global_batch_size $GBS --num_layers $NLS --hidden_size $HS --num_attention_heads $NAH \
 --DDP_impl local ${EXTRA:-} $COMMON"
. "$(dirname "$0")/SBATCH.sh"
