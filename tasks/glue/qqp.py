"""QQP TSV reader (tasks/glue/qqp.py:14-88 surface)."""
from epfl_megatron_b200 import print_rank_0
from data_utils import clean_text
from .data import GLUEAbstractDataset

LABELS = [0, 1]


class QQPDataset(GLUEAbstractDataset):
    def __init__(self, name, datapaths, tokenizer, max_seq_length, test_label=0):
        self.test_label = test_label
        super().__init__("QQP", name, datapaths, tokenizer, max_seq_length)

    def process_samples_from_single_path(self, filename):
        print_rank_0(f" > Processing {filename} ...")
        samples, is_test, skipped = [], False, 0
        with open(filename, encoding="utf-8") as f:
            for n, line in enumerate(f):
                row = line.strip().split("\t")
                if n == 0:
                    is_test = len(row) == 3             # id, question1, question2
                    continue
                if is_test:
                    if len(row) != 3:
                        skipped += 1
                        continue
                    uid, a, b, label = int(row[0]), row[1], row[2], self.test_label
                else:
                    if len(row) != 6:
                        skipped += 1
                        continue
                    uid, a, b, label = int(row[0]), row[3], row[4], int(row[5])
                a, b = clean_text(a.strip()), clean_text(b.strip())
                if not a or not b:
                    skipped += 1
                    continue
                assert label in LABELS and uid >= 0
                samples.append({"text_a": a, "text_b": b, "label": label, "uid": uid})
        print_rank_0(f" >> processed {len(samples)} samples ({skipped} malformed rows skipped).")
        return samples
