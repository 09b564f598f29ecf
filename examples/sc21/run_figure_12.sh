#!/bin/bash
# Interleaved vs plain 1F1B: bash run_figure_12.sh <yes|no> <GBS: 12|24|36|48|60>   (96 layers, h 12288, TP 8 x PP 12)
set -euo pipefail
. "$(dirname "$0")/CONFIG.sh"
INTERLEAVED=${1:-yes}; GBS=${2:-12}
TP=8; PP=12; MBS=1; NLS=96; HS=12288; NAH=96; NNODES=12
[ "$INTERLEAVED" = yes ] && EXTRA="--num_layers_per_virtual_pipeline_stage 2"
export NNODES
export OPTIONS="--tensor_model_parallel_size $TP --pipeline_model_parallel_size $PP --micro_batch_size $MBS \
 --global_batch_size $GBS --num_layers $NLS --hidden_size $HS --num_attention_heads $NAH \
 --DDP_impl local ${EXTRA:-} $COMMON"
. "$(dirname "$0")/SBATCH.sh"
