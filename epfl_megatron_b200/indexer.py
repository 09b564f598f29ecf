"""Embed every evidence block with the context tower of a biencoder and store the vectors
(megatron/indexer.py:17-123 surface). Each data-parallel rank embeds its slice and dumps a
shard; rank 0 merges them into ``--embedding_path``."""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import global_vars
from .checkpointing import load_biencoder_checkpoint
from .data.biencoder_dataset_utils import get_one_epoch_dataloader
from .data.orqa_wiki_dataset import get_open_retrieval_batch, get_open_retrieval_wiki_dataset
from .data.realm_index import OpenRetreivalDataStore, detach
from .models.biencoder_model import get_model_provider
from .parallel import state as mpu
from .training import get_model
from .utils import print_rank_0, unwrap_to


class IndexBuilder:
    def __init__(self):
        args = global_vars.get_args()
        self.model = self.dataloader = self.dataset = self.evidence_embedder_obj = None
        self.biencoder_shared_query_context_model = args.biencoder_shared_query_context_model
        # without --ict_load the towers must come from a BERT checkpoint
        self.using_realm_chkpt = args.ict_load is None
        self.log_interval = args.indexer_log_interval
        self.batch_size = args.indexer_batch_size
        self.load_attributes()
        self.is_main_builder = mpu.get_data_parallel_rank() == 0
        self.num_total_builders = mpu.get_data_parallel_world_size()
        self.iteration = self.total_processed = 0

    def load_attributes(self):
        """Context tower only (shared-tower models load the shared weights)."""
        only_context = not self.biencoder_shared_query_context_model
        model = get_model(get_model_provider(
            only_context_model=only_context,
            biencoder_shared_query_context_model=self.biencoder_shared_query_context_model))
        self.model = load_biencoder_checkpoint(model, only_context_model=only_context)
        assert len(self.model) == 1
        self.model[0].eval()
        self.dataset = get_open_retrieval_wiki_dataset()
        self.dataloader = iter(get_one_epoch_dataloader(self.dataset, self.batch_size))
        self.evidence_embedder_obj = OpenRetreivalDataStore(load_from_path=False)

    def track_and_report_progress(self, batch_size):
        self.iteration += 1
        self.total_processed += batch_size * self.num_total_builders
        if self.is_main_builder and self.iteration % self.log_interval == 0:
            print(f"Batch {self.iteration:10d} | Total {self.total_processed:10d}", flush=True)

    def build_and_save_index(self):
        model = unwrap_to(self.model[0], "embed_text")
        model.eval()
        while True:
            try:
                row_id, tokens, mask, types, _ = get_open_retrieval_batch(self.dataloader)
            except (StopIteration, IndexError):
                break
            assert mask.dtype == torch.bool
            with torch.no_grad():
                logits = model.embed_text(model.context_model, tokens, mask, types)
            self.evidence_embedder_obj.add_block_data(detach(row_id), detach(logits))
            self.track_and_report_progress(batch_size=len(row_id))
        self.evidence_embedder_obj.save_shard()
        if dist.is_initialized():
            dist.barrier()
        del self.model
        if self.is_main_builder:
            self.evidence_embedder_obj.merge_shards_and_save()
            assert len(self.evidence_embedder_obj.embed_data) == len(self.dataset)
        self.evidence_embedder_obj.clear()
        if dist.is_initialized():
            dist.barrier()
        print_rank_0("index built")
