#!/bin/bash
# Weak scaling of pipeline parallelism: bash run_figure_11.sh <PP: 1|2|4|8> <GBS: 8|128>   (3 layers per stage, h 20480, TP 8)
set -euo pipefail
. "$(dirname "$0")/CONFIG.sh"
PP=${1:?pipeline-parallel size}; GBS=${2:-8}
TP=8; MBS=1; NLS=$((3 * PP)); HS=20480; NAH=128; NNODES=$PP
export NNODES
export OPTIONS="--tensor_model_parallel_size $TP --pipeline_model_parallel_size $PP --micro_batch_size $MBS \
 --global_batch_size $GBS --num_layers $NLS --hidden_size $HS --num_attention_heads $NAH \
 --DDP_impl local ${EXTRA:-} $COMMON"
. "$(dirname "$0")/SBATCH.sh"
