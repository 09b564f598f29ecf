#!/bin/bash
# Token-level F1 of the generated responses against the references written by data_processing.sh.
set -euo pipefail
CKPT=${CHECKPOINT_PATH:?model checkpoint}
VOCAB=${VOCAB_PATH:?gpt2-vocab.json}
MERGES=${MERGE_PATH:?gpt2-merges.txt}
NGPU=${NGPU:-1}
MODEL="--num_layers 24 --hidden_size 1024 --num_attention_heads 16 --seq_length 2048 --max_position_embeddings 2048 \
 --micro_batch_size 1 --tokenizer_type GPT2BPETokenizer --vocab_file $VOCAB --merge_file $MERGES --load $CKPT --fp16"
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 6000"
$RUN tasks/msdp/main.py --task MSDP-EVAL-F1 --guess_file ${GUESS_FILE:?generations} \
  --answer_file ${ANSWER_FILE:?*_resp_ref.txt} $MODEL
