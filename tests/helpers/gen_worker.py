"""Subprocess body of tests/test_text_generation.py: a tiny random Llama on CPU/Gloo, cached
greedy decoding vs full re-computation, sampling, beam search and the REST server."""
import json
import os
import sys
import threading
import urllib.request

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    import torch
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    import finetune
    from epfl_megatron_b200 import get_args
    from epfl_megatron_b200.initialize import initialize_megatron
    from epfl_megatron_b200.models import LlamaModel
    from epfl_megatron_b200.text_generation import (beam_search_and_post_process, generate,
                                                    generate_and_post_process)
    from epfl_megatron_b200.text_generation_server import MegatronServer
    from epfl_megatron_b200.training import get_model
    from epfl_megatron_b200.utils import get_ltor_masks_and_position_ids

    argv = ("--model_name llama2 --num_layers 2 --hidden_size 64 --num_attention_heads 4 "
            "--num_attention_heads_kv 2 --ffn_hidden_size 128 --seq_length 64 --max_position_embeddings 64 "
            "--micro_batch_size 1 --global_batch_size 1 --use_rms_norm --glu_activation swiglu "
            "--no_tie_embed_logits --position_embedding_type rotary --hidden_dropout 0.0 "
            "--attention_dropout 0.0 --tokenizer_type NullTokenizer --vocab_size 96 "
            "--make_vocab_size_divisible_by 32 --distributed_backend gloo --no_bias_gelu_fusion "
            "--use_cpu_initialization --train_iters 1 --lr 1e-4").split()
    initialize_megatron(finetune.extra_args, {}, args_list=argv)
    args = get_args()

    def provider(pre_process=True, post_process=True):
        return LlamaModel(num_tokentypes=0, parallel_output=False, pre_process=pre_process,
                          post_process=post_process, version=2)
    model = get_model(provider, wrap_with_ddp=False)[0]
    model.eval()
    out = {}

    prompts = ["5 6 7 8 9", "11 12 13"]
    toks, lens, lp = generate(model, prompts=prompts, tokens_to_generate=8, top_k_sampling=1,
                              return_output_log_probs=True, use_eod_token_for_early_termination=False)
    out["shape"] = list(toks.shape)
    # oracle: re-run the full prefix through the model without a cache for every new token
    ok = True
    with torch.no_grad():
        for b, n0 in enumerate(lens.tolist()):
            seq = toks[b:b + 1, :n0].clone()
            for t in range(n0, toks.size(1)):
                m, _, pos = get_ltor_masks_and_position_ids(seq, 95, False, False, False)
                logits = model(seq, pos, m)
                nxt = int(torch.argmax(logits[0, -1].float()))
                if nxt != int(toks[b, t]):
                    ok = False
                seq = torch.cat([seq, toks[b:b + 1, t:t + 1]], dim=1)
    out["greedy_matches_uncached"] = ok
    out["logprobs_finite"] = bool(torch.isfinite(lp).all()) and lp.shape[1] == toks.shape[1] - 1

    texts, segs, logprobs, _ = generate_and_post_process(model, prompts=prompts, tokens_to_generate=4,
                                                         top_p_sampling=0.9, temperature=0.8, random_seed=3,
                                                         return_output_log_probs=True)
    out["sampled_texts"] = len(texts) == 2 and all(t.startswith(p) for t, p in zip(texts, prompts))
    # scoring only
    _, _, sc = generate(model, prompts=prompts, tokens_to_generate=0, return_output_log_probs=True)
    out["score_shape_ok"] = list(sc.shape) == [2, 4]
    r = beam_search_and_post_process(model, prompts=["5 6 7"], tokens_to_generate=6, beam_size=3,
                                     stop_token=95, num_return_gen=2, length_penalty=1.0)
    out["beam_ok"] = len(r[0]) == 2 and r[2][0] >= r[2][1]

    srv = MegatronServer(model)
    port = int(os.environ.get("GEN_TEST_PORT", "5917"))
    th = threading.Thread(target=srv.run, args=("127.0.0.1",), kwargs={"port": port}, daemon=True)
    th.start()
    import time
    time.sleep(0.5)

    def put(payload):
        req = urllib.request.Request(f"http://127.0.0.1:{port}/api", data=json.dumps(payload).encode(),
                                     method="PUT", headers={"Content-Type": "application/json"})
        try:
            with urllib.request.urlopen(req, timeout=60) as resp:
                return resp.status, json.loads(resp.read().decode())
        except urllib.error.HTTPError as e:
            return e.code, json.loads(e.read().decode())
    st, body = put({"prompts": ["5 6 7"], "tokens_to_generate": 3, "top_k": 1, "logprobs": True})
    out["server_ok"] = st == 200 and body["text"][0].startswith("5 6 7") and len(body["logprobs"][0]) > 0
    st, body = put({"prompts": "not a list"})
    out["server_rejects"] = st == 400
    st, body = put({"prompts": ["5 6 7"], "tokens_to_generate": 4, "beam_width": 2, "stop_token": 95})
    out["server_beam"] = st == 200 and len(body["text"]) >= 1
    srv.shutdown()
    print("RESULT " + json.dumps(out))


if __name__ == "__main__":
    main()
