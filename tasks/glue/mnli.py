"""MNLI TSV reader (tasks/glue/mnli.py:14-71 surface)."""
from epfl_megatron_b200 import print_rank_0
from data_utils import clean_text
from .data import GLUEAbstractDataset

LABELS = {"contradiction": 0, "entailment": 1, "neutral": 2}


class MNLIDataset(GLUEAbstractDataset):
    def __init__(self, name, datapaths, tokenizer, max_seq_length, test_label="contradiction"):
        self.test_label = test_label
        super().__init__("MNLI", name, datapaths, tokenizer, max_seq_length)

    def process_samples_from_single_path(self, filename):
        print_rank_0(f" > Processing {filename} ...")
        samples, is_test = [], False
        with open(filename, encoding="utf-8") as f:
            for n, line in enumerate(f):
                row = line.strip().split("\t")
                if n == 0:
                    is_test = len(row) == 10            # the test split has no label columns
                    if is_test:
                        print_rank_0(f"   reading test data, label {self.test_label} will be assigned")
                    continue
                text_a, text_b = clean_text(row[8].strip()), clean_text(row[9].strip())
                label = self.test_label if is_test else row[-1].strip()
                assert text_a and text_b and label in LABELS
                samples.append({"text_a": text_a, "text_b": text_b, "label": LABELS[label], "uid": int(row[0])})
                if len(samples) % 50000 == 0:
                    print_rank_0(f"  > processed {len(samples)} so far ...")
        print_rank_0(f" >> processed {len(samples)} samples.")
        return samples
