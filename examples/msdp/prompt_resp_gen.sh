#!/bin/bash
# Stage 2: the language model writes the response given dialogue context and generated knowledge.
set -euo pipefail
CKPT=${CHECKPOINT_PATH:?model checkpoint}
VOCAB=${VOCAB_PATH:?gpt2-vocab.json}
MERGES=${MERGE_PATH:?gpt2-merges.txt}
NGPU=${NGPU:-1}
MODEL="--num_layers 24 --hidden_size 1024 --num_attention_heads 16 --seq_length 2048 --max_position_embeddings 2048 \
 --micro_batch_size 1 --tokenizer_type GPT2BPETokenizer --vocab_file $VOCAB --merge_file $MERGES --load $CKPT --fp16"
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 6000"
$RUN tasks/msdp/main.py --task MSDP-PROMPT --prompt_type response --num_prompt_examples 20 \
  --prompt_file ${PROMPT_PATH:?resp_prompts.txt} --sample_input_file ${INPUT_PATH:?output of prep_resp_gen.sh} \
  --sample_output_file ${OUTPUT_PATH:?generated responses} --out_seq_length 100 $MODEL
