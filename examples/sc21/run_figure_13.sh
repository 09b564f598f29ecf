#!/bin/bash
# Tensor vs pipeline parallelism on 64 GPUs: bash run_figure_13.sh <PP: 2|4|8|16|32> <GBS: 32|128>   (TP = 64 / PP)
set -euo pipefail
. "$(dirname "$0")/CONFIG.sh"
PP=${1:?pipeline-parallel size}; GBS=${2:-32}
TP=$((64 / PP)); MBS=1; NLS=32; HS=20480; NAH=128; NNODES=8
export NNODES
export OPTIONS="--tensor_model_parallel_size $TP --pipeline_model_parallel_size $PP --micro_batch_size $MBS \
 --global_batch_size $GBS --num_layers $NLS --hidden_size $HS --num_attention_heads $NAH \
 --DDP_impl local ${EXTRA:-} $COMMON"
. "$(dirname "$0")/SBATCH.sh"
