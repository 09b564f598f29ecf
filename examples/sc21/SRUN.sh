#!/bin/bash
export CUDA_DEVICE_MAX_CONNECTIONS=1
srun python pretrain_gpt.py $OPTIONS
