"""Base dataset of the GLUE tasks (tasks/glue/data.py:14-56 surface)."""
from abc import ABC, abstractmethod

from torch.utils.data import Dataset

from epfl_megatron_b200 import print_rank_0
from data_utils import build_sample, build_tokens_types_paddings_from_text


class GLUEAbstractDataset(ABC, Dataset):
    def __init__(self, task_name, dataset_name, datapaths, tokenizer, max_seq_length):
        self.task_name, self.dataset_name = task_name, dataset_name
        self.tokenizer, self.max_seq_length = tokenizer, max_seq_length
        print_rank_0(f" > building {task_name} dataset for {dataset_name}:")
        print_rank_0("   " + ", ".join(datapaths))
        self.samples = []
        for p in datapaths:
            self.samples.extend(self.process_samples_from_single_path(p))
        print_rank_0(f"  >> total number of samples: {len(self.samples)}")

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        raw = self.samples[idx]
        ids, types, paddings = build_tokens_types_paddings_from_text(raw["text_a"], raw["text_b"],
                                                                     self.tokenizer, self.max_seq_length)
        return build_sample(ids, types, paddings, raw["label"], raw["uid"])

    @abstractmethod
    def process_samples_from_single_path(self, datapath):
        """-> list of {text_a, text_b, label, uid}"""
