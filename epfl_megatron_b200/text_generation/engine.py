"""Inference engine: KV-cache forward step (with micro-batched pipelining), scoring, the
sampling loop and beam search.

Surface of megatron/text_generation/forward_step.py:17-204 and generation.py:20-429
(``InferenceParams``, ``ForwardStep``, ``score_and_return_on_first_stage``,
``generate_tokens_probs_and_return_on_first_stage``, ``beam_search_and_return_on_first_stage``).
Differences from the reference: the keys in the cache are stored already rotated (its RoPE +
cache combination rotates cached keys with the new tokens' table, transformer.py:494-501), the
`[1,1,s,s]` mask is built once per call, and the end-of-generation test is the only host
synchronisation per token.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import global_vars
from ..parallel import state as mpu
from ..utils import get_ltor_masks_and_position_ids, unwrap_model
from .beam_utils import BeamHypotheses
from .communication import (broadcast_from_last_pipeline_stage,
                            broadcast_from_last_to_first_pipeline_stage,
                            copy_from_last_to_first_pipeline_stage, recv_from_prev_pipeline_rank_,
                            send_to_next_pipeline_rank)
from .sampling import sample


class InferenceParams:
    """Per-request decoding state handed to every attention layer: cache geometry, the write
    offsets of the current call and the per-layer (key, value) buffers."""

    def __init__(self, max_batch_size, max_sequence_len):
        self.max_sequence_len = max_sequence_len
        self.max_batch_size = max_batch_size
        self.sequence_len_offset = 0
        self.batch_size_offset = 0
        self.key_value_memory_dict = {}

    def swap_key_value_dict(self, batch_idx):
        """Reorder the cached sequences (beam search keeps the surviving beams)."""
        if not self.key_value_memory_dict:
            raise ValueError("the KV cache is empty")
        for layer, (k, v) in self.key_value_memory_dict.items():
            assert len(batch_idx) == k.shape[1]
            self.key_value_memory_dict[layer] = (k[:, batch_idx], v[:, batch_idx])


def _model_device(model):
    return next(model.parameters()).device


class ForwardStep:
    """model(tokens, positions, mask) -> logits on the last stage, with the KV cache advanced.
    Large prompts are cut into micro-batches when the pipeline has several stages so that
    the stages overlap."""

    def __init__(self, model, max_batch_size, max_sequence_len):
        assert not isinstance(model, (list, tuple)), "interleaved schedules are not supported for inference"
        model.eval()
        self.model = model
        self.inference_params = InferenceParams(max_batch_size, max_sequence_len)
        args = global_vars.get_args()
        self.pipeline_size_larger_than_one = args.pipeline_model_parallel_size > 1
        self.pipelining_batch_x_seqlen = args.inference_batch_times_seqlen_threshold

    def __call__(self, tokens, position_ids, attention_mask):
        if self.pipeline_size_larger_than_one:
            b, s = tokens.size(0), tokens.size(1)
            if b * s >= self.pipelining_batch_x_seqlen:
                mbs = max(1, self.pipelining_batch_x_seqlen // s)
                return self._pipelined(tokens, position_ids, attention_mask, mbs)
        return self._plain(tokens, position_ids, attention_mask)

    def _stage(self, tokens, position_ids, attention_mask, recv_buffer=None):
        args = global_vars.get_args()
        b, s = tokens.size(0), tokens.size(1)
        if not mpu.is_pipeline_first_stage():
            if recv_buffer is None:
                dt = torch.float if args.fp32_residual_connection else args.params_dtype
                recv_buffer = torch.empty(s, b, args.hidden_size, dtype=dt, device=tokens.device)
            recv_from_prev_pipeline_rank_(recv_buffer)
            unwrap_model(self.model).set_input_tensor(recv_buffer)
        else:
            unwrap_model(self.model).set_input_tensor(None)
        out = self.model(tokens, position_ids, attention_mask, inference_params=self.inference_params)
        send_to_next_pipeline_rank(out)
        return out

    def _plain(self, tokens, position_ids, attention_mask):
        out = self._stage(tokens, position_ids, attention_mask)
        self.inference_params.sequence_len_offset += tokens.size(1)
        return out if mpu.is_pipeline_last_stage() else None

    def _pipelined(self, tokens, position_ids, attention_mask, micro_batch_size):
        b, s = tokens.size(0), tokens.size(1)
        logits = None
        for start in range(0, b, micro_batch_size):
            end = min(b, start + micro_batch_size)
            self.inference_params.batch_size_offset = start
            out = self._stage(tokens[start:end], position_ids[start:end], attention_mask)
            if mpu.is_pipeline_last_stage():
                if logits is None:
                    logits = torch.empty(b, s, out.size(-1), dtype=out.dtype, device=out.device)
                logits[start:end] = out
        self.inference_params.batch_size_offset = 0
        self.inference_params.sequence_len_offset += s
        return logits


def _build_attention_mask_and_position_ids(tokens):
    """Causal mask [1,1,s,s] (True = masked out) and positions [b,s] for left-to-right decoding."""
    tokenizer = global_vars.get_tokenizer()
    mask, _, position_ids = get_ltor_masks_and_position_ids(
        tokens, tokenizer.eod, reset_position_ids=False, reset_attention_mask=False, eod_mask_loss=False)
    return mask, position_ids


def _gather_full_vocab(logits):
    """Generation models are built with parallel_output=False, so logits are already full."""
    return logits


def score_and_return_on_first_stage(model, tokens, lengths):
    """Log-probabilities of the given tokens under the model (no generation).
    Returns tokens, lengths, output_log_probs [b, s-1] (valid on the first stage)."""
    args = global_vars.get_args()
    b, s = tokens.size(0), tokens.size(1)
    if s > args.max_position_embeddings:
        raise ValueError("Length of prompt + tokens_to_generate longer than allowed")
    if s * b > args.max_tokens_to_oom:
        raise ValueError(f"Too many tokens. {s * b} is greater than {args.max_tokens_to_oom}")
    forward_step = ForwardStep(model, b, s)
    out_size = (b, s - 1)
    log_probs = None
    with torch.no_grad():
        mask, position_ids = _build_attention_mask_and_position_ids(tokens)
        logits = forward_step(tokens, position_ids, mask)
        if mpu.is_pipeline_last_stage():
            lp = F.log_softmax(logits.float(), dim=2)
            log_probs = torch.gather(lp, 2, tokens[:, 1:].unsqueeze(2)).squeeze(2)
    log_probs = broadcast_from_last_to_first_pipeline_stage(out_size, torch.float32, log_probs)
    return tokens, lengths, log_probs


def generate_tokens_probs_and_return_on_first_stage(
        model, tokens, lengths, return_output_log_probs=False, top_k=0, top_p=0.0, top_p_decay=0.0,
        top_p_bound=0.0, temperature=1.0, use_eod_token_for_early_termination=True,
        stop_on_double_eol=False, stop_on_eol=False, prevent_newline_after_colon=True):
    """The sampling loop. tokens [b, max_len] holds the prompts padded with EOD; lengths [b] the
    prompt lengths. Decoding starts at the shortest prompt; rows whose prompt is longer keep their
    prompt token until they are reached. Returns (tokens, total lengths, log-probs or None),
    valid on the first pipeline stage."""
    args = global_vars.get_args()
    tokenizer = global_vars.get_tokenizer()
    dev = tokens.device
    b, max_len = tokens.size(0), tokens.size(1)
    min_prompt = int(lengths.min().item())
    if max_len > args.max_position_embeddings:
        raise ValueError("Length of prompt + tokens_to_generate longer than allowed")
    if max_len * b > args.max_tokens_to_oom:
        raise ValueError(f"Too many tokens. {max_len * b} is greater than {args.max_tokens_to_oom}")
    forward_step = ForwardStep(model, b, max_len)
    termination_id = getattr(args, "eos_id", None)
    if termination_id is None:
        termination_id = tokenizer.eod
    colon = newline = None
    if prevent_newline_after_colon:
        try:
            colon, newline = tokenizer.tokenize(":")[0], tokenizer.tokenize("\n")[0]
        except Exception:  # noqa: BLE001 - tokenizers without these symbols
            prevent_newline_after_colon = False
    last = mpu.is_pipeline_last_stage()
    out_lp = torch.empty(b, max_len - 1, dtype=torch.float32, device=dev) if (last and return_output_log_probs) else None
    gen_lengths = torch.full((b,), max_len, dtype=torch.int64, device=dev) if last else None
    is_done = torch.zeros(b, dtype=torch.bool, device=dev)
    context_length = min_prompt
    with torch.no_grad():
        mask, position_ids = _build_attention_mask_and_position_ids(tokens)
        prev = 0
        for context_length in range(min_prompt, max_len):
            logits = forward_step(tokens[:, prev:context_length], position_ids[:, prev:context_length],
                                  mask[..., prev:context_length, :context_length])
            new_sample = None
            if last:
                logits = logits.float()
                if prevent_newline_after_colon:
                    logits[tokens[:, context_length - 1] == colon, -1, newline] = -1e10
                new_sample = sample(logits[:, -1, :], top_k=top_k, top_p=top_p, temperature=temperature,
                                    vocab_size=tokenizer.vocab_size)
                if top_p > 0.0 and top_p_decay > 0.0:
                    top_p = top_p * top_p_decay
                    if top_p_bound > 0.0:
                        top_p = max(top_p, top_p_bound)
                started = lengths <= context_length
                tokens[started, context_length] = new_sample[started]
                if return_output_log_probs:
                    lp = F.log_softmax(logits, dim=2)
                    idx = tokens[:, prev + 1:context_length + 1].unsqueeze(2)
                    out_lp[:, prev:context_length] = torch.gather(lp, 2, idx).squeeze(2)
            copy_from_last_to_first_pipeline_stage(b, torch.int64, tokens[:, context_length])
            prev = context_length
            done = None
            if last:
                if stop_on_double_eol:
                    hit = ((new_sample == 628) | ((new_sample == 198) & (tokens[:, context_length - 1] == 198))) & started
                elif stop_on_eol:
                    hit = ((new_sample == 628) | (new_sample == 198)) & started
                else:
                    hit = (new_sample == termination_id) & started
                just = hit & ~is_done
                gen_lengths[just] = context_length + 1
                is_done = is_done | hit
                done = torch.all(is_done).to(torch.uint8).view(1)
            done = broadcast_from_last_pipeline_stage(1, torch.uint8, tensor=done)
            if use_eod_token_for_early_termination and bool(done.item()):
                break
    tokens = tokens[:, :context_length + 1]
    if last and return_output_log_probs:
        out_lp = out_lp[:, :context_length]
    gen_lengths = broadcast_from_last_to_first_pipeline_stage(b, torch.int64, gen_lengths)
    if return_output_log_probs:
        out_lp = broadcast_from_last_to_first_pipeline_stage((b, context_length), torch.float32, out_lp)
    return tokens, gen_lengths, out_lp


def beam_search_and_return_on_first_stage(model, tokens, lengths, beam_size, stop_token, num_return_gen,
                                          length_penalty, prevent_newline_after_colon=True):
    """Beam search for ONE prompt. Returns (tokens [num_return_gen, len], scores) on the first stage."""
    args = global_vars.get_args()
    tokenizer = global_vars.get_tokenizer()
    assert tokens.size(0) == 1, "beam search takes a single prompt"
    dev = tokens.device
    prompt_length = int(lengths.item())
    final_len = tokens.size(1)
    if prompt_length >= final_len:
        raise ValueError("context length + tokens_to_generate too large")
    forward_step = ForwardStep(model, beam_size, final_len)
    hyps = BeamHypotheses(beam_size, length_penalty)
    best_batches = None
    done = torch.zeros(1, dtype=torch.uint8, device=dev)
    scores = torch.zeros(beam_size, dtype=torch.float32, device=dev).unsqueeze(1)
    scores_size = None
    colon = newline = None
    if prevent_newline_after_colon:
        try:
            colon, newline = tokenizer.tokenize(":")[0], tokenizer.tokenize("\n")[0]
        except Exception:  # noqa: BLE001
            prevent_newline_after_colon = False
    last = mpu.is_pipeline_last_stage()
    with torch.no_grad():
        tokens = tokens.repeat(beam_size, 1)
        mask, position_ids = _build_attention_mask_and_position_ids(tokens)
        prev = 0
        context_length = prompt_length
        for context_length in range(prompt_length, final_len):
            logits = forward_step(tokens[:, prev:context_length], position_ids[:, prev:context_length],
                                  mask[..., prev:context_length, :context_length])
            if last:
                logits = logits.float()
                if prevent_newline_after_colon:
                    logits[tokens[:, context_length - 1] == colon, -1, newline] = -1e10
                vocab = logits.size(2)
                lp = F.log_softmax(logits, dim=2)
                new_scores = lp[:, -1, :] + scores
                if context_length == prompt_length:     # all beams are identical: use beam 0
                    sorted_scores, idx = torch.sort(new_scores[0, :], descending=True)
                else:
                    sorted_scores, idx = torch.sort(new_scores.view(-1), descending=True)
                best_beam_ids = torch.div(idx[:2 * beam_size], vocab, rounding_mode="floor")
                best_words = idx[:2 * beam_size] % vocab
                best_scores = sorted_scores[:2 * beam_size]
                nxt = []
                for rank_, (tok, sc, beam) in enumerate(zip(best_words.tolist(), best_scores.tolist(),
                                                            best_beam_ids.tolist())):
                    if tok == stop_token:
                        if rank_ >= beam_size:      # a finished hypothesis outside the top beams is dropped
                            continue
                        hyps.add(tokens[beam].clone(), sc, context_length + 1 - prompt_length)
                    else:
                        nxt.append((tok, sc, beam))
                    if len(nxt) == beam_size:
                        break
                if hyps.is_done(float(best_scores.max().item()), context_length + 1 - prompt_length):
                    done = torch.ones(1, dtype=torch.uint8, device=dev)
                best_batches = tokens.new_tensor([x[2] for x in nxt])
                tokens = tokens[best_batches, :]
                tokens[:, context_length] = tokens.new_tensor([x[0] for x in nxt])
                scores = scores.new_tensor([x[1] for x in nxt]).unsqueeze(1)
            done = broadcast_from_last_pipeline_stage(1, torch.uint8, done)
            if bool(done.item()):
                break
            # every stage needs the surviving beams and their caches
            copy_from_last_to_first_pipeline_stage(tokens.size(), torch.int64, tokens)
            best_batches = broadcast_from_last_pipeline_stage(beam_size, torch.int64, best_batches)
            forward_step.inference_params.swap_key_value_dict(best_batches)
            prev = context_length
        if last:
            if not bool(done.item()):               # ran out of length: close the open beams
                for i in range(beam_size):
                    hyps.add(tokens[i].clone(), float(scores[i].item()), context_length + 1 - prompt_length)
            ranked = sorted(hyps.beams, key=lambda x: x[0], reverse=True)
            n = min(num_return_gen, len(ranked))
            out_scores = torch.stack([torch.tensor(ranked[i][0], device=dev) for i in range(n)])
            out_tokens = torch.stack([ranked[i][1] for i in range(n)])
            scores_size = torch.tensor(out_scores.shape, dtype=torch.int64, device=dev)
            tokens_size = torch.tensor(out_tokens.shape, dtype=torch.int64, device=dev)
        else:
            out_scores = out_tokens = tokens_size = None
        scores_size = broadcast_from_last_pipeline_stage(1, torch.int64, scores_size)
        tokens_size = broadcast_from_last_pipeline_stage(2, torch.int64, tokens_size)
        out_scores = broadcast_from_last_to_first_pipeline_stage(tuple(scores_size.tolist()), torch.float32, out_scores)
        out_tokens = broadcast_from_last_to_first_pipeline_stage(tuple(tokens_size.tolist()), torch.int64, out_tokens)
    return out_tokens, out_scores
