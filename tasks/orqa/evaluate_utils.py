"""Retrieval accuracy of the query tower against a pre-built evidence index
(tasks/orqa/evaluate_utils.py:16-180 surface): embed the questions, search the MIPS index on
one rank per node, share the top-k ids, then score answer presence."""
import torch
import torch.distributed as dist

from epfl_megatron_b200 import get_args, print_rank_0
from epfl_megatron_b200.checkpointing import load_biencoder_checkpoint
from epfl_megatron_b200.data.orqa_wiki_dataset import get_open_retrieval_wiki_dataset
from epfl_megatron_b200.data.realm_index import FaissMIPSIndex, OpenRetreivalDataStore
from epfl_megatron_b200.models.biencoder_model import get_model_provider
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.training import get_model
from epfl_megatron_b200.utils import unwrap_to
from orqa.unsupervised.nq import get_nq_dataset, get_one_epoch_nq_dataloader, process_nq_batch
from orqa.unsupervised.qa_utils import calculate_matches


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class ORQAEvaluator:
    def __init__(self):
        args = get_args()
        self.embedding_size = args.hidden_size
        self.faiss_use_gpu = args.faiss_use_gpu
        self.evidence_embedder_obj = None
        self.evidence_dataset = None
        self.mips_index = None
        self.eval_dataset = None
        only_query = not args.biencoder_shared_query_context_model
        model = get_model(get_model_provider(
            only_query_model=only_query,
            biencoder_shared_query_context_model=args.biencoder_shared_query_context_model))
        self.model = load_biencoder_checkpoint(model, only_query_model=only_query)
        assert len(self.model) == 1
        self.model[0].eval()
        self.faiss_wrapper()

    def get_evidence_embedding(self):
        self.evidence_embedder_obj = OpenRetreivalDataStore(load_from_path=True)

    def get_evidence_dataset(self):
        self.evidence_dataset = get_open_retrieval_wiki_dataset()

    def faiss_wrapper(self):
        """One rank per node (local rank 0) holds the index; the others wait."""
        args = get_args()
        if args.local_rank in (0, None):
            self.get_evidence_embedding()
            assert self.evidence_embedder_obj is not None
            self.mips_index = FaissMIPSIndex(embed_size=self.embedding_size,
                                             embed_data=self.evidence_embedder_obj, use_gpu=self.faiss_use_gpu)
        self.get_evidence_dataset()
        if dist.is_initialized():
            dist.barrier()

    def generate_query_vectors(self, qa_data, split):
        self.eval_dataset = get_nq_dataset(qa_data, split)
        loader = get_one_epoch_nq_dataloader(self.eval_dataset)
        vectors, reference = [], []
        for batch in loader:
            tokens, mask, types, _, refs = process_nq_batch(batch)
            assert len(self.model) == 1
            um = unwrap_to(self.model[0], "embed_text")
            with torch.no_grad():
                q = um.embed_text(um.query_model, tokens.to(_dev()), mask.to(_dev()), types.to(_dev()))
            reference.extend(refs)
            vectors.extend(q.split(1, dim=0))
        return torch.cat(vectors, dim=0).contiguous(), reference

    def evaluate(self, qa_data, split):
        args = get_args()
        query, reference = self.generate_query_vectors(qa_data, split)
        world = dist.get_world_size(mpu.get_data_parallel_group()) if dist.is_initialized() else 1
        if world > 1:
            parts = [torch.empty_like(query) for _ in range(world)]
            dist.all_gather(parts, query, group=mpu.get_data_parallel_group())
            all_query = torch.cat(parts, dim=0)
        else:
            all_query = query
        topk = args.faiss_topk_retrievals
        n = all_query.shape[0]
        distance = torch.zeros(n, topk, dtype=torch.float32)
        ids = torch.zeros(n, topk, dtype=torch.int64)
        if self.mips_index is not None:
            d, i = self.mips_index.search_mips_index(all_query, top_k=topk, reconstruct=False)
            distance[:, :d.shape[1]] = torch.from_numpy(d)
            ids[:, :i.shape[1]] = torch.from_numpy(i)
        if world > 1:
            distance, ids = distance.to(_dev()), ids.to(_dev())
            src = (dist.get_rank() // max(torch.cuda.device_count(), 1)) * max(torch.cuda.device_count(), 1) \
                if torch.cuda.is_available() else 0
            dist.broadcast(distance, src=src)
            dist.broadcast(ids, src=src)
            rank = dist.get_rank(mpu.get_data_parallel_group())
            per = n // world
            distance, ids = distance[rank * per:(rank + 1) * per], ids[rank * per:(rank + 1) * per]
        top_ids_and_scores = [(ids[k].tolist(), distance[k].tolist()) for k in range(ids.shape[0])]
        stats = calculate_matches(self.evidence_dataset.id2text, reference, top_ids_and_scores,
                                  workers_num=args.num_workers, match_type=args.faiss_match)
        hits = stats.top_k_hits
        print_rank_0(f"{split} SET RESULTS")
        print_rank_0(f"topk-{topk} documents hits {hits}")
        acc = [v / len(top_ids_and_scores) for v in hits]
        for k in args.retriever_report_topk_accuracies:
            if k <= len(acc):
                print_rank_0(f"top-{k}: {acc[k - 1] * 100:.2f}")
        return acc
