#!/bin/bash
# Stage 0: turn the raw Wizard-of-Wikipedia / Wizard-of-Internet dumps into the tab-separated format of
# tasks/msdp and select the prompts for both generation stages.
set -euo pipefail
DIR=${MSDP_DIR:?directory with the raw datasets}
WOW=$DIR/wizard_of_wikipedia
WOI=$DIR/wizard_of_internet
ENCODER=${DPR_ENCODER:-}        # optional sentence encoder checkpoint for prompt retrieval

for split in train test_random_split test_topic_split; do
  python tasks/msdp/preprocessing.py --func process_wow_dataset --raw_file $WOW/$split.json \
    --processed_file $WOW/${split}_processed.txt --knwl_ref_file $WOW/${split}_knwl_ref.txt \
    --resp_ref_file $WOW/${split}_resp_ref.txt
done
for split in train test; do
  python tasks/msdp/preprocessing.py --func process_woi_dataset --raw_file $WOI/$split.jsonl \
    --processed_file $WOI/${split}_processed.txt --knwl_ref_file $WOI/${split}_knwl_ref.txt \
    --resp_ref_file $WOI/${split}_resp_ref.txt
done
# knowledge-generation prompts: nearest training dialogues per test sample
python tasks/msdp/preprocessing.py --func get_knwl_gen_prompts --test_file $WOW/test_random_split_processed.txt \
  --train_file $WOW/train_processed.txt ${ENCODER:+--model_file $ENCODER} \
  --processed_file $WOW/knwl_prompts_seen.jsonl --data_type wow_seen
python tasks/msdp/preprocessing.py --func get_knwl_gen_prompts --test_file $WOW/test_topic_split_processed.txt \
  --train_file $WOW/train_processed.txt ${ENCODER:+--model_file $ENCODER} \
  --processed_file $WOW/knwl_prompts_unseen.jsonl --data_type wow_unseen
python tasks/msdp/preprocessing.py --func get_knwl_gen_prompts --test_file $WOI/test_processed.txt \
  --train_file $WOI/train_processed.txt ${ENCODER:+--model_file $ENCODER} \
  --processed_file $WOI/knwl_prompts.jsonl --data_type woi
# response-generation prompts: a fixed random sample of training turns
python tasks/msdp/preprocessing.py --func get_resp_gen_prompts --train_file $WOW/train_processed.txt \
  --processed_file $WOW/resp_prompts.txt --seed 1234
