"""Public generation API: ``generate_and_post_process``, ``generate``, ``beam_search_and_post_process``,
``beam_search`` with the keyword arguments and return values of megatron/text_generation/api.py:19-201.

Only rank 0 has to pass real arguments. Every call first synchronises its scalar options over all ranks
as ONE float vector (``_REQUEST`` / ``_BEAM_REQUEST`` name the fields and their types), then the prompts
as token tensors; after that all ranks run the same engine call and the first pipeline stage holds the
result.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..parallel import state as mpu
from . import engine
from .communication import broadcast_float_list
from .tokenization import detokenize_generations, tokenize_prompts

# scalar options of a sampling request, in wire order
_REQUEST = (("tokens_to_generate", int), ("return_output_log_probs", bool), ("top_k_sampling", int),
            ("top_p_sampling", float), ("top_p_decay", float), ("top_p_bound", float), ("temperature", float),
            ("add_BOS", bool), ("use_eod_token_for_early_termination", bool), ("stop_on_double_eol", bool),
            ("stop_on_eol", bool), ("prevent_newline_after_colon", bool), ("random_seed", int))
_BEAM_REQUEST = (("tokens_to_generate", int), ("beam_size", int), ("add_BOS", bool), ("stop_token", int),
                 ("num_return_gen", int), ("length_penalty", float), ("prevent_newline_after_colon", bool))


def _agree(fields, given: dict) -> dict:
    """Rank 0's values of ``fields`` on every rank, restored to their types."""
    wire = broadcast_float_list(len(fields), float_list=[float(given[name]) for name, _ in fields])
    return {name: kind(wire[i].item()) for i, (name, kind) in enumerate(fields)}


def _on_rank_zero() -> bool:
    return not dist.is_initialized() or dist.get_rank() == 0


def generate(model, prompts=None, tokens_to_generate=0, return_output_log_probs=False, top_k_sampling=0,
             top_p_sampling=0.0, top_p_decay=0.0, top_p_bound=0.0, temperature=1.0, add_BOS=False,
             use_eod_token_for_early_termination=True, stop_on_double_eol=False, stop_on_eol=False,
             prevent_newline_after_colon=False, random_seed=-1):
    """-> (tokens [b, len], lengths [b], log_probs or None) on the first pipeline stage.
    ``tokens_to_generate == 0`` scores the prompts instead of extending them."""
    req = _agree(_REQUEST, locals())
    if req["random_seed"] != -1:
        torch.random.manual_seed(req["random_seed"])
    if _on_rank_zero():
        assert prompts is not None
    context, lengths = tokenize_prompts(prompts=prompts, tokens_to_generate=req["tokens_to_generate"],
                                        add_BOS=req["add_BOS"])
    if req["tokens_to_generate"] == 0:
        return engine.score_and_return_on_first_stage(model, context, lengths)
    passthrough = ("return_output_log_probs", "top_p_decay", "top_p_bound", "temperature",
                   "use_eod_token_for_early_termination", "stop_on_double_eol", "stop_on_eol",
                   "prevent_newline_after_colon")
    return engine.generate_tokens_probs_and_return_on_first_stage(
        model, context, lengths, top_k=req["top_k_sampling"], top_p=req["top_p_sampling"],
        **{name: req[name] for name in passthrough})


def generate_and_post_process(model, prompts=None, tokens_to_generate=0, return_output_log_probs=False,
                              top_k_sampling=0, top_p_sampling=0.0, top_p_decay=0.0, top_p_bound=0.0,
                              temperature=1.0, add_BOS=False, use_eod_token_for_early_termination=True,
                              stop_on_double_eol=False, stop_on_eol=False,
                              prevent_newline_after_colon=False, random_seed=-1):
    """``generate`` + detokenisation on the host: (texts, segments, log_probs, tokens) on the first
    pipeline stage, None elsewhere. Log-probabilities are cut to the generated length of each sample."""
    options = dict(locals())
    options.pop("model")
    tokens, lengths, log_probs = generate(model, **options)
    if not mpu.is_pipeline_first_stage():
        return None
    tokens, texts, segments = detokenize_generations(tokens, lengths, True)
    if return_output_log_probs:
        log_probs = [row[:len(seg) - 1] for row, seg in zip(log_probs.cpu().numpy().tolist(), segments)]
    return texts, segments, log_probs, tokens


def beam_search(model, prompts=None, tokens_to_generate=0, beam_size=0, add_BOS=False, stop_token=50256,
                num_return_gen=1, length_penalty=1, prevent_newline_after_colon=False):
    """-> (tokens [num_return_gen, len], scores) on the first pipeline stage."""
    req = _agree(_BEAM_REQUEST, locals())
    context, lengths = tokenize_prompts(prompts=prompts, tokens_to_generate=req["tokens_to_generate"],
                                        add_BOS=req["add_BOS"])
    return engine.beam_search_and_return_on_first_stage(
        model, context, lengths, req["beam_size"], stop_token=req["stop_token"],
        num_return_gen=req["num_return_gen"], length_penalty=req["length_penalty"],
        prevent_newline_after_colon=req["prevent_newline_after_colon"])


def beam_search_and_post_process(model, prompts=None, tokens_to_generate=0, beam_size=0, add_BOS=False,
                                 stop_token=50256, num_return_gen=1, length_penalty=1,
                                 prevent_newline_after_colon=False):
    """``beam_search`` + detokenisation: (texts, segments, scores) on the first pipeline stage, None elsewhere."""
    options = dict(locals())
    options.pop("model")
    tokens, scores = beam_search(model, **options)
    if not mpu.is_pipeline_first_stage():
        return None
    lengths = torch.full((tokens.size(0),), tokens.size(1), dtype=torch.int64, device=tokens.device)
    _, texts, segments = detokenize_generations(tokens, lengths, True)
    return texts, segments, scores.cpu().numpy().tolist()
