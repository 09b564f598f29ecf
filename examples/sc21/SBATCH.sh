#!/bin/bash
mkdir -p "$LOG_DIR"
sbatch --nodes="$NNODES" --ntasks-per-node=8 --exclusive --output="$LOG_DIR/%j.log" "$(dirname "$0")/SRUN.sh"
