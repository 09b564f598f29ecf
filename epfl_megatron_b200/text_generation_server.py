"""REST front end for generation: ``PUT /api`` with a JSON body, the request format of
megatron/text_generation_server.py:31-241 (prompts, tokens_to_generate, logprobs, temperature,
top_k, top_p, top_p_decay, top_p_bound, add_BOS, stop_on_double_eol, stop_on_eol,
prevent_newline_after_colon, random_seed, no_log, beam_width, stop_token, length_penalty).

The reference uses Flask; this one is built on the standard library's http.server, so the
serving rank has no extra dependency. The first pipeline stage of TP rank 0 serves; every other
rank sits in ``serve_worker_loop`` and joins each request after an opcode broadcast.
"""
from __future__ import annotations

import datetime
import json
import os
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import torch
import torch.distributed as dist

from .text_generation import beam_search_and_post_process, generate_and_post_process
from .text_generation.communication import _dev

GENERATE_NUM = 0
BEAM_NUM = 1
_LOCK = threading.Lock()


def _send_opcode(code: int):
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(torch.tensor([code], dtype=torch.int64, device=_dev()), 0)


def serve_worker_loop(model):
    """Ranks other than 0: wait for an opcode, then take part in the collective generation."""
    while True:
        choice = torch.zeros(1, dtype=torch.int64, device=_dev())
        dist.broadcast(choice, 0)
        if int(choice[0].item()) == GENERATE_NUM:
            generate_and_post_process(model)
        elif int(choice[0].item()) == BEAM_NUM:
            beam_search_and_post_process(model)
        else:
            return


def validate_request(req: dict):
    """-> (params dict, None) or (None, error message). Mirrors the reference's checks."""
    if "prompts" not in req:
        return None, "prompts argument required"
    if "max_len" in req:
        return None, "max_len is no longer used.  Replace with tokens_to_generate"
    if "sentences" in req:
        return None, "sentences is no longer used.  Replace with prompts"
    prompts = req["prompts"]
    if not isinstance(prompts, list) or not all(isinstance(p, str) for p in prompts):
        return None, "prompts is not a list of strings"
    if len(prompts) == 0:
        return None, "prompts is empty"
    if len(prompts) > 128:
        return None, "Maximum number of prompts is 128"
    p = dict(prompts=prompts, tokens_to_generate=64, logprobs=False, temperature=1.0, top_k=0, top_p=0.0,
             top_p_decay=0.0, top_p_bound=0.0, add_BOS=False, stop_on_double_eol=False, stop_on_eol=False,
             prevent_newline_after_colon=False, random_seed=-1, no_log=False, beam_width=None,
             stop_token=50256, length_penalty=1.0)

    def is_int(x):
        return isinstance(x, int) and not isinstance(x, bool)

    def is_num(x):
        return isinstance(x, (int, float)) and not isinstance(x, bool)

    if "tokens_to_generate" in req:
        v = req["tokens_to_generate"]
        if not is_int(v) or v < 0:
            return None, "tokens_to_generate must be an integer greater than or equal to 0"
        p["tokens_to_generate"] = v
    if "logprobs" in req:
        if not isinstance(req["logprobs"], bool):
            return None, "logprobs must be a boolean value"
        p["logprobs"] = req["logprobs"]
    if p["tokens_to_generate"] == 0 and not p["logprobs"]:
        return None, "tokens_to_generate=0 implies logprobs should be True"
    if "temperature" in req:
        v = req["temperature"]
        if not is_num(v) or not (0.0 < v <= 100.0):
            return None, "temperature must be a positive number less than or equal to 100.0"
        p["temperature"] = float(v)
    if "top_k" in req:
        v = req["top_k"]
        if not is_int(v) or not (0 <= v <= 1000):
            return None, "top_k must be an integer equal to or greater than 0 and less than or equal to 1000"
        p["top_k"] = v
    if "top_p" in req:
        v = req["top_p"]
        if not is_num(v) or not (0.0 <= v <= 1.0):
            return None, "top_p must be a positive float less than or equal to 1.0"
        if v > 0.0 and p["top_k"] > 0:
            return None, "cannot set both top-k and top-p samplings."
        p["top_p"] = float(v)
    if "top_p_decay" in req:
        v = req["top_p_decay"]
        if not is_num(v) or not (0.0 <= v <= 1.0):
            return None, "top_p_decay must be a positive float less than or equal to 1.0"
        if p["top_p"] == 0.0:
            return None, "top_p_decay cannot be set without top_p"
        p["top_p_decay"] = float(v)
    if "top_p_bound" in req:
        v = req["top_p_bound"]
        if not is_num(v):
            return None, "top_p_bound must be a positive float less than or equal to top_p"
        if p["top_p"] == 0.0:
            return None, "top_p_bound cannot be set without top_p"
        if not (0.0 < v <= p["top_p"]):
            return None, "top_p_bound must be greater than 0 and less than top_p"
        p["top_p_bound"] = float(v)
    for key in ("add_BOS", "stop_on_double_eol", "stop_on_eol", "prevent_newline_after_colon", "no_log"):
        if key in req:
            if not isinstance(req[key], bool):
                return None, f"{key} must be a boolean value"
            p[key] = req[key]
    if any(len(s) == 0 for s in prompts) and not p["add_BOS"]:
        return None, "Empty prompts require add_BOS=true"
    if "random_seed" in req:
        v = req["random_seed"]
        if not is_int(v):
            return None, "random_seed must be integer"
        if v < 0:
            return None, "random_seed must be a positive integer"
        p["random_seed"] = v
    if "beam_width" in req:
        v = req["beam_width"]
        if not is_int(v):
            return None, "beam_width must be integer"
        if v < 1:
            return None, "beam_width must be an integer > 1"
        if len(prompts) > 1:
            return None, "When doing beam_search, batch size must be 1"
        p["beam_width"] = v
    if "stop_token" in req:
        if not is_int(req["stop_token"]):
            return None, "stop_token must be an integer"
        p["stop_token"] = req["stop_token"]
    if "length_penalty" in req:
        if not is_num(req["length_penalty"]):
            return None, "length_penalty must be a float"
        p["length_penalty"] = float(req["length_penalty"])
    return p, None


def handle_request(model, req: dict):
    """-> (status, body). Runs the request on all ranks (this rank drives)."""
    p, err = validate_request(req)
    if err:
        return 400, err
    with _LOCK:   # one generation at a time: the ranks move in lock step
        if not p["no_log"]:
            print("request IP/time:", datetime.datetime.now(), json.dumps(req), flush=True)
        try:
            if p["beam_width"] is not None:
                _send_opcode(BEAM_NUM)
                texts, segments, scores = beam_search_and_post_process(
                    model, prompts=p["prompts"], tokens_to_generate=p["tokens_to_generate"],
                    beam_size=p["beam_width"], add_BOS=p["add_BOS"], stop_token=p["stop_token"],
                    num_return_gen=p["beam_width"], length_penalty=p["length_penalty"],
                    prevent_newline_after_colon=p["prevent_newline_after_colon"])
                return 200, {"text": texts, "segments": segments, "scores": scores}
            _send_opcode(GENERATE_NUM)
            texts, segments, logprobs, _ = generate_and_post_process(
                model, prompts=p["prompts"], tokens_to_generate=p["tokens_to_generate"],
                return_output_log_probs=p["logprobs"], top_k_sampling=p["top_k"], top_p_sampling=p["top_p"],
                top_p_decay=p["top_p_decay"], top_p_bound=p["top_p_bound"], temperature=p["temperature"],
                add_BOS=p["add_BOS"], use_eod_token_for_early_termination=True,
                stop_on_double_eol=p["stop_on_double_eol"], stop_on_eol=p["stop_on_eol"],
                prevent_newline_after_colon=p["prevent_newline_after_colon"], random_seed=p["random_seed"])
            return 200, {"text": texts, "segments": segments, "logprobs": logprobs}
        except ValueError as e:
            return 400, e.args[0] if e.args else str(e)


class MegatronServer:
    """``MegatronServer(model).run("0.0.0.0", port=5000)`` -- blocks and serves PUT /api and the
    static test page at ``/``."""

    def __init__(self, model):
        self.model = model
        self.httpd = None

    def _handler(self):
        model = self.model
        static = os.path.join(os.path.dirname(__file__), "static", "index.html")

        class Handler(BaseHTTPRequestHandler):
            def _reply(self, status, body, ctype="application/json"):
                data = body if isinstance(body, bytes) else (
                    json.dumps(body) if not isinstance(body, str) else json.dumps(body)).encode()
                self.send_response(status)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(data)))
                self.end_headers()
                self.wfile.write(data)

            def do_PUT(self):   # noqa: N802
                if self.path.rstrip("/") != "/api":
                    return self._reply(404, "not found")
                try:
                    n = int(self.headers.get("Content-Length", "0"))
                    req = json.loads(self.rfile.read(n) or b"{}")
                except (ValueError, json.JSONDecodeError):
                    return self._reply(400, "request body is not JSON")
                status, body = handle_request(model, req)
                self._reply(status, body)

            do_POST = do_PUT    # noqa: N815

            def do_GET(self):   # noqa: N802
                if self.path in ("/", "/index.html") and os.path.exists(static):
                    with open(static, "rb") as f:
                        return self._reply(200, f.read(), "text/html")
                self._reply(404, "not found")

            def log_message(self, fmt, *a):   # quiet
                pass
        return Handler

    def run(self, url, port=5000):
        self.httpd = ThreadingHTTPServer((url, port), self._handler())
        try:
            self.httpd.serve_forever()
        finally:
            self.httpd.server_close()

    def shutdown(self):
        if self.httpd is not None:
            self.httpd.shutdown()
