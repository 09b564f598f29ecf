#!/bin/bash
set -euo pipefail
: "${EVIDENCE:?}" "${EMBEDDINGS:?}" "${VOCAB:?}" "${CHECKPOINT_PATH:?}" "${QA_DEV:?}"
torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tasks/main.py --task RETRIEVER-EVAL \
  --tokenizer_type BertWordPieceLowerCase --num_layers 12 --hidden_size 768 --num_attention_heads 12 \
  --micro_batch_size 128 --seq_length 512 --max_position_embeddings 512 --load "$CHECKPOINT_PATH" \
  --evidence_data_path "$EVIDENCE" --embedding_path "$EMBEDDINGS" --retriever_seq_length 256 \
  --vocab_file "$VOCAB" --qa_data_dev "$QA_DEV" --faiss_use_gpu --retriever_report_topk_accuracies 1 5 20 100 --bf16
