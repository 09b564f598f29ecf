#!/bin/bash
# Activation recompute on / off: bash run_figure_17.sh <yes|no> <GBS: 1|2|4|8|16|32|64|128|256>   (TP 8 x PP 16)
set -euo pipefail
. "$(dirname "$0")/CONFIG.sh"
RECOMPUTE=${1:-yes}; GBS=${2:-1}
TP=8; PP=16; MBS=1; NLS=80; HS=12288; NAH=96; NNODES=16
# CONFIG.sh turns full recompute on for the whole sweep; this figure needs it switchable
[ "$RECOMPUTE" = yes ] || COMMON=${COMMON/--recompute_granularity full --recompute_method uniform/}
export NNODES
export OPTIONS="--tensor_model_parallel_size $TP --pipeline_model_parallel_size $PP --micro_batch_size $MBS \
 --global_batch_size $GBS --num_layers $NLS --hidden_size $HS --num_attention_heads $NAH \
 --DDP_impl local ${EXTRA:-} $COMMON"
. "$(dirname "$0")/SBATCH.sh"
