#!/bin/bash
# Between the stages: put the generated knowledge next to each test turn.
set -euo pipefail
python tasks/msdp/preprocessing.py --func prepare_input --test_file ${TEST_FILE:?test_*_processed.txt} \
  --knwl_gen_file ${KNWL_GEN_FILE:?output of prompt_knwl_gen.sh} --processed_file ${PROCESSED_FILE:?response-stage input}
