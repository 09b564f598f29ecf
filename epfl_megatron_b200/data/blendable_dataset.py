"""Weighted mixture of datasets (megatron/data/blendable_dataset.py:12-53 behaviour)."""
import time

import numpy as np
import torch

from ..utils import print_rank_0
from . import helpers


class BlendableDataset(torch.utils.data.Dataset):
    def __init__(self, datasets, weights):
        self.datasets = datasets
        n = len(datasets)
        assert n == len(weights) and n < 255
        self.size = sum(len(d) for d in datasets)
        w = np.array(weights, dtype=np.float64)
        s = np.sum(w)
        assert s > 0.0
        w /= s
        t0 = time.time()
        self.dataset_index = np.zeros(self.size, dtype=np.uint8)
        self.dataset_sample_index = np.zeros(self.size, dtype=np.int64)
        rank0 = not torch.distributed.is_initialized() or torch.distributed.get_rank() == 0
        helpers.build_blending_indices(self.dataset_index, self.dataset_sample_index, w, n, self.size, rank0)
        print_rank_0("> elapsed time for building blendable dataset indices: {:.2f} (sec)".format(time.time() - t0))

    def __len__(self):
        return self.size

    def __getitem__(self, idx):
        return self.datasets[self.dataset_index[idx]][self.dataset_sample_index[idx]]
