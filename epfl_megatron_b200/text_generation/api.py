"""Public generation API (megatron/text_generation/api.py:19-201): same function names,
keyword arguments and return values."""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..parallel import state as mpu
from .communication import broadcast_float_list
from .engine import (beam_search_and_return_on_first_stage,
                     generate_tokens_probs_and_return_on_first_stage, score_and_return_on_first_stage)
from .tokenization import detokenize_generations, tokenize_prompts


def generate_and_post_process(model, prompts=None, tokens_to_generate=0, return_output_log_probs=False,
                              top_k_sampling=0, top_p_sampling=0.0, top_p_decay=0.0, top_p_bound=0.0,
                              temperature=1.0, add_BOS=False, use_eod_token_for_early_termination=True,
                              stop_on_double_eol=False, stop_on_eol=False,
                              prevent_newline_after_colon=False, random_seed=-1):
    """Generate, detokenise, move to the host. Returns (texts, segments, log_probs, tokens) on the
    first pipeline stage and None elsewhere."""
    tokens, lengths, output_log_probs = generate(
        model, prompts=prompts, tokens_to_generate=tokens_to_generate,
        return_output_log_probs=return_output_log_probs, top_k_sampling=top_k_sampling,
        top_p_sampling=top_p_sampling, top_p_decay=top_p_decay, top_p_bound=top_p_bound,
        temperature=temperature, add_BOS=add_BOS,
        use_eod_token_for_early_termination=use_eod_token_for_early_termination,
        stop_on_double_eol=stop_on_double_eol, stop_on_eol=stop_on_eol,
        prevent_newline_after_colon=prevent_newline_after_colon, random_seed=random_seed)
    if not mpu.is_pipeline_first_stage():
        return None
    tokens, texts, segments = detokenize_generations(tokens, lengths, True)
    if return_output_log_probs:
        output_log_probs = output_log_probs.cpu().numpy().tolist()
        for i, (lp, seg) in enumerate(zip(output_log_probs, segments)):
            output_log_probs[i] = lp[:len(seg) - 1]
    return texts, segments, output_log_probs, tokens


def generate(model, prompts=None, tokens_to_generate=0, return_output_log_probs=False, top_k_sampling=0,
             top_p_sampling=0.0, top_p_decay=0.0, top_p_bound=0.0, temperature=1.0, add_BOS=False,
             use_eod_token_for_early_termination=True, stop_on_double_eol=False, stop_on_eol=False,
             prevent_newline_after_colon=False, random_seed=-1):
    """-> (tokens [b, len], lengths [b], log_probs or None). Only rank 0 needs real arguments:
    they are broadcast as one float vector, the prompts as token tensors."""
    values = [tokens_to_generate, return_output_log_probs, top_k_sampling, top_p_sampling, top_p_decay,
              top_p_bound, temperature, add_BOS, use_eod_token_for_early_termination, stop_on_double_eol,
              stop_on_eol, prevent_newline_after_colon, random_seed]
    v = broadcast_float_list(len(values), float_list=[float(x) for x in values])
    tokens_to_generate = int(v[0].item())
    return_output_log_probs = bool(v[1].item())
    top_k_sampling = int(v[2].item())
    top_p_sampling, top_p_decay, top_p_bound, temperature = (v[i].item() for i in (3, 4, 5, 6))
    add_BOS = bool(v[7].item())
    use_eod_token_for_early_termination = bool(v[8].item())
    stop_on_double_eol, stop_on_eol = bool(v[9].item()), bool(v[10].item())
    prevent_newline_after_colon = bool(v[11].item())
    random_seed = int(v[12].item())
    if random_seed != -1:
        torch.random.manual_seed(random_seed)
    if not dist.is_initialized() or dist.get_rank() == 0:
        assert prompts is not None
    context_tokens, context_lengths = tokenize_prompts(prompts=prompts, tokens_to_generate=tokens_to_generate,
                                                       add_BOS=add_BOS)
    if tokens_to_generate == 0:
        return score_and_return_on_first_stage(model, context_tokens, context_lengths)
    return generate_tokens_probs_and_return_on_first_stage(
        model, context_tokens, context_lengths, return_output_log_probs=return_output_log_probs,
        top_k=top_k_sampling, top_p=top_p_sampling, top_p_decay=top_p_decay, top_p_bound=top_p_bound,
        temperature=temperature, use_eod_token_for_early_termination=use_eod_token_for_early_termination,
        stop_on_double_eol=stop_on_double_eol, stop_on_eol=stop_on_eol,
        prevent_newline_after_colon=prevent_newline_after_colon)


def beam_search_and_post_process(model, prompts=None, tokens_to_generate=0, beam_size=0, add_BOS=False,
                                 stop_token=50256, num_return_gen=1, length_penalty=1,
                                 prevent_newline_after_colon=False):
    """-> (texts, segments, scores) on the first pipeline stage, None elsewhere."""
    tokens, scores = beam_search(model, prompts=prompts, tokens_to_generate=tokens_to_generate,
                                 beam_size=beam_size, add_BOS=add_BOS, stop_token=stop_token,
                                 num_return_gen=num_return_gen, length_penalty=length_penalty,
                                 prevent_newline_after_colon=prevent_newline_after_colon)
    if not mpu.is_pipeline_first_stage():
        return None
    lengths = torch.full((tokens.size(0),), tokens.size(1), dtype=torch.int64, device=tokens.device)
    tokens, texts, segments = detokenize_generations(tokens, lengths, True)
    return texts, segments, scores.cpu().numpy().tolist()


def beam_search(model, prompts=None, tokens_to_generate=0, beam_size=0, add_BOS=False, stop_token=50256,
                num_return_gen=1, length_penalty=1, prevent_newline_after_colon=False):
    values = [tokens_to_generate, beam_size, add_BOS, stop_token, num_return_gen, length_penalty,
              prevent_newline_after_colon]
    v = broadcast_float_list(len(values), float_list=[float(x) for x in values])
    tokens_to_generate, beam_size = int(v[0].item()), int(v[1].item())
    add_BOS, stop_token = bool(v[2].item()), int(v[3].item())
    num_return_gen, length_penalty = int(v[4].item()), v[5].item()
    prevent_newline_after_colon = bool(v[6].item())
    context_tokens, context_lengths = tokenize_prompts(prompts=prompts, tokens_to_generate=tokens_to_generate,
                                                       add_BOS=add_BOS)
    return beam_search_and_return_on_first_stage(
        model, context_tokens, context_lengths, beam_size, stop_token=stop_token,
        num_return_gen=num_return_gen, length_penalty=length_penalty,
        prevent_newline_after_colon=prevent_newline_after_colon)
