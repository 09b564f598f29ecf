"""RACE reading-comprehension reader: every question becomes NUM_CHOICES (context, question+option)
pairs (tasks/race/data.py:17-135 surface)."""
import glob
import json
import os
import time

from torch.utils.data import Dataset

from epfl_megatron_b200 import print_rank_0
from data_utils import build_sample, build_tokens_types_paddings_from_ids, clean_text

NUM_CHOICES = 4
MAX_QA_LENGTH = 128


class RaceDataset(Dataset):
    def __init__(self, dataset_name, datapaths, tokenizer, max_seq_length, max_qa_length=MAX_QA_LENGTH):
        self.dataset_name = dataset_name
        print_rank_0(f" > building RACE dataset for {dataset_name}:")
        print_rank_0("   " + ", ".join(datapaths))
        self.samples = []
        for p in datapaths:
            self.samples.extend(process_single_datapath(p, tokenizer, max_qa_length, max_seq_length))
        print_rank_0(f"  >> total number of samples: {len(self.samples)}")
        self.sample_multiplier = NUM_CHOICES      # each sample carries NUM_CHOICES sequences

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        return self.samples[idx]


def process_single_datapath(datapath, tokenizer, max_qa_length, max_seq_length):
    print_rank_0(f"   > working on {datapath}")
    t0 = time.time()
    samples, uid = [], 0
    for fname in sorted(glob.glob(os.path.join(datapath, "*.txt"))):
        with open(fname, encoding="utf-8") as f:
            for line in f:
                data = json.loads(line)
                context_ids = tokenizer.tokenize(clean_text(data["article"]))
                for q, opts, ans in zip(data["questions"], data["options"], data["answers"]):
                    assert len(opts) == NUM_CHOICES
                    label = ord(ans) - ord("A")
                    assert 0 <= label < NUM_CHOICES
                    ids_l, types_l, pad_l = [], [], []
                    for opt in opts:
                        # cloze questions carry a blank that the option fills
                        qa = q.replace("_", opt) if "_" in q else " ".join([q, opt])
                        qa_ids = tokenizer.tokenize(clean_text(qa))[:max_qa_length]
                        ids, types, paddings = build_tokens_types_paddings_from_ids(
                            qa_ids, context_ids, max_seq_length, tokenizer.cls, tokenizer.sep, tokenizer.pad)
                        ids_l.append(ids)
                        types_l.append(types)
                        pad_l.append(paddings)
                    samples.append(build_sample(ids_l, types_l, pad_l, label, uid))
                    uid += 1
    print_rank_0(f"    > processed {len(samples)} samples in {time.time() - t0:.2f} seconds")
    return samples
