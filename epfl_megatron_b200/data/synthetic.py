"""Synthetic token dataset (``--synthetic_data``): deterministic pseudo-random token
streams of the requested shape, for benchmarks and tests on machines without a corpus.
No counterpart in the reference (its tokenizer / dataset need files or network)."""
import numpy as np
import torch


class SyntheticGPTDataset(torch.utils.data.Dataset):
    def __init__(self, num_samples, seq_length, vocab_size, seed=1234):
        self.num_samples = int(num_samples)
        self.seq_length, self.vocab_size, self.seed = seq_length, vocab_size, seed

    def __len__(self):
        return self.num_samples

    def __getitem__(self, idx):
        rng = np.random.default_rng(self.seed + int(idx))
        return {"text": rng.integers(0, self.vocab_size, size=self.seq_length + 1, dtype=np.int64)}


def build_synthetic_datasets(args, train_val_test_num_samples):
    vocab = args.padded_vocab_size
    return tuple(SyntheticGPTDataset(max(n, 1), args.seq_length, vocab, args.seed + i * 7919)
                 for i, n in enumerate(train_val_test_num_samples))
