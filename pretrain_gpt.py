"""``pretrain_gpt.py`` entry point. The reference's tools and launch scripts still call it
(tools/checkpoint_loader_megatron.py:91, examples/pretrain_gpt_distributed_with_mp.sh:19)
although the file itself is gone from that tree; here it is finetune.py with
``--model_name gpt`` as the default."""
import sys

import finetune
from finetune import (extra_args, forward_step, get_batch, loss_func, model_provider,  # noqa: F401
                      train_valid_test_datasets_provider)

if __name__ == "__main__":
    if not any(a.startswith("--model_name") for a in sys.argv):
        sys.argv += ["--model_name", "gpt"]
    finetune.main()
