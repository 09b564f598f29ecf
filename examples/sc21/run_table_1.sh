#!/bin/bash
# bash examples/sc21/run_table_1.sh <1.7B|3.6B|7.5B|18B|39B|76B|145B|310B|530B|1T>
set -euo pipefail
. "$(dirname "$0")/CONFIG.sh"
case ${1:?model size} in
  1.7B) TP=1; PP=1; MBS=16; GBS=512; NLS=24; HS=2304; NAH=24; NNODES=4;;
  3.6B) TP=2; PP=1; MBS=16; GBS=512; NLS=30; HS=3072; NAH=32; NNODES=8;;
  7.5B) TP=4; PP=1; MBS=16; GBS=512; NLS=36; HS=4096; NAH=32; NNODES=16;;
  18B) TP=8; PP=1; MBS=8; GBS=1024; NLS=40; HS=6144; NAH=48; NNODES=32;;
  39B) TP=8; PP=2; MBS=4; GBS=1536; NLS=48; HS=8192; NAH=64; NNODES=64;;
  76B) TP=8; PP=4; MBS=2; GBS=1792; NLS=60; HS=10240; NAH=80; NNODES=128;;
  145B) TP=8; PP=8; MBS=2; GBS=2304; NLS=80; HS=12288; NAH=96; NNODES=192;;
  310B) TP=8; PP=16; MBS=1; GBS=2160; NLS=96; HS=16384; NAH=128; NNODES=240;;
  530B) TP=8; PP=35; MBS=1; GBS=2520; NLS=105; HS=20480; NAH=128; NNODES=315;;
  1T) TP=8; PP=64; MBS=1; GBS=3072; NLS=128; HS=25600; NAH=160; NNODES=384;;
  *) echo "unknown size $1"; exit 1;;
esac
export NNODES
export OPTIONS="--tensor_model_parallel_size $TP --pipeline_model_parallel_size $PP --micro_batch_size $MBS \
 --global_batch_size $GBS --num_layers $NLS --hidden_size $HS --num_attention_heads $NAH \
 --sequence_parallel --use_distributed_optimizer $COMMON"
. "$(dirname "$0")/SBATCH.sh"
