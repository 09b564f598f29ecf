"""Block-embedding stores for retrieval: a dict-backed shard store that every rank dumps and
rank 0 merges, and a (optionally GPU) FAISS inner-product index
(megatron/data/realm_index.py:15-224 surface). FAISS is optional; without it a dense torch
matmul search is used."""
from __future__ import annotations

import itertools
import os
import pickle
import shutil

import numpy as np
import torch

from .. import global_vars
from ..parallel import state as mpu


def detach(t):
    return t.detach().cpu().numpy()


class OpenRetreivalDataStore:
    """embed_data: {row id -> embedding (fp16 numpy)}; serialised as a pickle."""

    def __init__(self, embedding_path=None, load_from_path=True, rank=None):
        self.embed_data = {}
        if embedding_path is None:
            embedding_path = global_vars.get_args().embedding_path
            rank = global_vars.get_args().rank
        self.embedding_path, self.rank = embedding_path, rank
        if load_from_path:
            self.load_from_file()
        self.temp_dir_name = os.path.splitext(self.embedding_path)[0] + "_tmp"

    def state(self):
        return {"embed_data": self.embed_data}

    def clear(self):
        """Drop the embeddings (they are no longer needed once an index was built)."""
        self.embed_data = {}

    def load_from_file(self):
        if not mpu.model_parallel_is_initialized() or mpu.get_data_parallel_rank() == 0:
            print("\n> Unpickling BlockData", flush=True)
        with open(self.embedding_path, "rb") as f:
            self.embed_data = pickle.load(f)["embed_data"]
        if not mpu.model_parallel_is_initialized() or mpu.get_data_parallel_rank() == 0:
            print(">> Finished unpickling BlockData\n", flush=True)

    def add_block_data(self, row_id, block_embeds, allow_overwrite=False):
        for idx, emb in zip(row_id, block_embeds):
            if not allow_overwrite and idx in self.embed_data:
                raise ValueError("Unexpectedly tried to overwrite block data")
            self.embed_data[idx] = np.float16(emb)

    def save_shard(self):
        os.makedirs(self.temp_dir_name, exist_ok=True)
        with open(f"{self.temp_dir_name}/{self.rank}.pkl", "wb") as f:
            pickle.dump(self.state(), f)

    def merge_shards_and_save(self):
        """Rank 0 only: fold every shard into one pickle and remove the temp dir. The merging rank's
        own shard is already in memory when it embedded blocks itself (megatron/data/realm_index.py:86-115)."""
        seen = set()
        own_in_memory = bool(self.embed_data)
        for fname in os.listdir(self.temp_dir_name):
            rank = int(os.path.splitext(fname)[0])
            seen.add(rank)
            if own_in_memory and rank == self.rank:
                continue
            with open(f"{self.temp_dir_name}/{fname}", "rb") as f:
                data = pickle.load(f)
                before, n = len(self.embed_data), len(data["embed_data"])
                self.embed_data.update(data["embed_data"])
                assert len(self.embed_data) == before + n, "shards overlap"
        assert len(seen) == len(os.listdir(self.temp_dir_name))
        with open(self.embedding_path, "wb") as f:
            pickle.dump(self.state(), f)
        shutil.rmtree(self.temp_dir_name, ignore_errors=True)
        print(f"Finished merging {len(seen)} shards for a total of {len(self.embed_data)} embeds", flush=True)


class FaissMIPSIndex:
    """Maximum-inner-product search over block embeddings."""

    def __init__(self, embed_size, embed_data=None, use_gpu=False):
        self.embed_size, self.embed_data, self.use_gpu = embed_size, embed_data, use_gpu
        self.mips_index = None
        self._dense = None          # (ids, matrix) fallback when faiss is missing
        self._set_mips_index()

    def _set_mips_index(self):
        try:
            import faiss
        except ImportError:
            self.mips_index = None
            if not mpu.model_parallel_is_initialized() or mpu.get_data_parallel_rank() == 0:
                print(">> faiss not available: dense torch search is used", flush=True)
            if self.embed_data:
                self.add_embed_data(self.embed_data)
            return
        cpu_index = faiss.IndexFlatIP(self.embed_size)
        if self.use_gpu:
            cfg = faiss.GpuMultipleClonerOptions()
            cfg.useFloat16, cfg.shard = True, True
            self.mips_index = faiss.IndexIDMap(faiss.index_cpu_to_all_gpus(cpu_index, co=cfg))
        else:
            self.mips_index = faiss.IndexIDMap(cpu_index)
        if self.embed_data:
            self.add_embed_data(self.embed_data)

    def reset_index(self):
        del self.mips_index
        self._dense = None
        if self.embed_data is not None:
            emb = self.embed_data.embed_data
            if emb and not isinstance(next(iter(emb.values())), np.ndarray):
                self.embed_data.embed_data = {k: np.float16(v) for k, v in emb.items()}
        self._set_mips_index()

    def update_index(self):
        self.embed_data.load_from_file()
        self.reset_index()

    def add_embed_data(self, all_embed_data):
        ids, embeds = zip(*all_embed_data.embed_data.items())
        embeds = np.float32(np.array(embeds))
        all_embed_data.clear()
        if self.mips_index is not None:
            self.mips_index.add_with_ids(embeds, np.array(ids, dtype=np.int64))
        else:
            self._dense = (torch.tensor(ids, dtype=torch.int64), torch.from_numpy(embeds))

    def search_mips_index(self, query_embeds, top_k, reconstruct=True):
        """-> (scores, ids) [n, top_k]; with reconstruct=True returns the embeddings instead."""
        q = np.float32(detach(query_embeds)) if torch.is_tensor(query_embeds) else np.float32(query_embeds)
        if self.mips_index is not None:
            if reconstruct:
                return self.mips_index.search_and_reconstruct(q, top_k)
            return self.mips_index.search(q, top_k)
        ids, mat = self._dense
        scores = torch.from_numpy(q) @ mat.t()
        top, idx = torch.topk(scores, min(top_k, mat.shape[0]), dim=1)
        if reconstruct:
            return mat[idx].numpy()
        return top.numpy(), ids[idx].numpy()
