"""Unit tests of the small runtime pieces (the reference's tests/test_basic.py, test_utils.py,
test_activations.py family): schedules, micro-batch calculators, loss scaler,
left-to-right masks, GLU activations, RoPE, timers, tokenizer padding / SentencePiece specials,
checkpoint file naming."""
import math
import types

import pytest
import torch

from epfl_megatron_b200.microbatches import ConstantNumMicroBatches, RampupBatchsizeNumMicroBatches
from epfl_megatron_b200.optimizer.grad_scaler import ConstantGradScaler, DynamicGradScaler
from epfl_megatron_b200.optimizer_param_scheduler import OptimizerParamScheduler
from epfl_megatron_b200.utils import get_ltor_masks_and_position_ids


class _Opt:
    def __init__(self):
        self.param_groups = [{"lr": 0.0, "weight_decay": 0.0, "wd_mult": 1.0, "lr_mult": 1.0},
                             {"lr": 0.0, "weight_decay": 0.0, "wd_mult": 0.0, "lr_mult": 2.0}]


def _sched(style, **kw):
    d = dict(max_lr=1.0, min_lr=0.1, lr_warmup_steps=10, lr_decay_steps=110, lr_decay_style=style,
             start_wd=0.0, end_wd=0.2, wd_incr_steps=100, wd_incr_style="linear")
    d.update(kw)
    opt = _Opt()
    return OptimizerParamScheduler(opt, **d), opt


def test_lr_schedules_in_samples():
    s, opt = _sched("cosine")
    assert opt.param_groups[0]["lr"] == 0.0
    s.step(5)
    assert opt.param_groups[0]["lr"] == pytest.approx(0.5)            # linear warm-up
    assert opt.param_groups[1]["lr"] == pytest.approx(1.0)            # lr multiplier group
    s.step(5)
    assert s.get_lr() == pytest.approx(1.0)
    s.step(50)                                                        # halfway through the decay
    assert s.get_lr() == pytest.approx(0.1 + 0.9 * 0.5 * (math.cos(math.pi * 0.5) + 1))
    assert opt.param_groups[0]["weight_decay"] == pytest.approx(0.2 * 0.6)
    assert opt.param_groups[1]["weight_decay"] == 0.0                 # no-decay group
    s.step(1000)
    assert s.get_lr() == pytest.approx(0.1) and s.get_wd() == pytest.approx(0.2)
    lin, _ = _sched("linear")
    lin.step(60)
    assert lin.get_lr() == pytest.approx(0.1 + 0.9 * 0.5)
    const, _ = _sched("constant")
    const.step(10 ** 6)
    assert const.get_lr() == 1.0
    isr, _ = _sched("inverse-square-root")
    isr.step(40)
    assert isr.get_lr() == pytest.approx(1.0 * 10 ** 0.5 / 40 ** 0.5)
    cosw, _ = _sched("cosine", wd_incr_style="cosine")
    cosw.step(50)
    assert cosw.get_wd() == pytest.approx(0.1)


def test_scheduler_checkpoint_reconciliation():
    a, _ = _sched("cosine")
    a.step(37)
    sd = a.state_dict()
    b, opt = _sched("cosine")
    b.load_state_dict(sd)
    assert b.num_steps == 37 and opt.param_groups[0]["lr"] == pytest.approx(a.get_lr())
    strict, _ = _sched("cosine", max_lr=2.0, use_checkpoint_opt_param_scheduler=False)
    with pytest.raises(AssertionError):
        strict.load_state_dict(sd)
    over, _ = _sched("cosine", max_lr=2.0, use_checkpoint_opt_param_scheduler=False, override_opt_param_scheduler=True)
    over.load_state_dict(sd)
    assert over.max_lr == 2.0
    legacy = {"start_lr": 1.0, "min_lr": 0.1, "warmup_iter": 10, "end_iter": 110, "decay_style": "cosine", "num_iters": 5}
    c, _ = _sched("cosine")
    c.load_state_dict(legacy)
    assert c.num_steps == 5


def test_num_microbatch_calculators():
    c = ConstantNumMicroBatches(64, 2, 4)
    assert c.get() == 8 and c.get_current_global_batch_size() == 64
    with pytest.raises(AssertionError):
        ConstantNumMicroBatches(30, 4, 2)
    r = RampupBatchsizeNumMicroBatches(16, 16, 300, 64, 2, 4)
    seen = []
    for consumed in (0, 99, 100, 199, 200, 300, 301, 10 ** 6):
        r.update(consumed, True)
        seen.append((r.get_current_global_batch_size(), r.get()))
    assert seen == [(16, 2), (16, 2), (32, 4), (32, 4), (48, 6), (64, 8), (64, 8), (64, 8)]
    bad = RampupBatchsizeNumMicroBatches(12, 4, 100, 64, 2, 4)     # 12 is not a multiple of mbs * dp
    with pytest.raises(AssertionError):
        bad.update(0, True)


def test_dynamic_grad_scaler_hysteresis():
    s = DynamicGradScaler(2.0 ** 10, 1.0, 2.0, 0.5, growth_interval=3, hysteresis=2)
    s.update(True)
    assert float(s.scale) == 2.0 ** 10                       # first overflow is tolerated
    s.update(True)
    assert float(s.scale) == 2.0 ** 9
    for _ in range(3):
        s.update(False)
    assert float(s.scale) == 2.0 ** 10 and float(s.inv_scale) == 2.0 ** -10
    s.update(True)
    assert float(s.scale) == 2.0 ** 10                       # hysteresis was re-armed by the growth
    floor = DynamicGradScaler(2.0, 1.0, 2.0, 0.5, 10, 1)
    for _ in range(5):
        floor.update(True)
    assert float(floor.scale) == 1.0
    t = DynamicGradScaler(4.0, 1.0, 2.0, 0.5, 10, 1)
    t.load_state_dict(s.state_dict())
    assert float(t.scale) == float(s.scale)
    c = ConstantGradScaler(128.0)
    c.update(True)
    assert float(c.scale) == 128.0


def test_ltor_masks_and_position_resets():
    eod = 0
    data = torch.tensor([[5, 6, eod, 7, 8, 9, eod, 4], [1, 2, 3, 4, 5, 6, 7, 8]])
    mask, loss_mask, pos = get_ltor_masks_and_position_ids(data, eod, True, True, True)
    assert pos[0].tolist() == [0, 1, 2, 0, 1, 2, 3, 0] and pos[1].tolist() == list(range(8))
    assert loss_mask[0].tolist() == [1, 1, 0, 1, 1, 1, 0, 1]
    vis = ~mask                       # True = may attend
    assert vis.shape == (2, 1, 8, 8)
    assert vis[0, 0, 4, 3] and not vis[0, 0, 4, 2] and not vis[0, 0, 3, 0] and not vis[0, 0, 2, 3]
    assert vis[1, 0, 7].all()
    mask2, loss2, pos2 = get_ltor_masks_and_position_ids(data, eod, False, False, False)
    assert mask2.shape == (1, 1, 8, 8) and loss2.sum() == 16 and pos2[0].tolist() == list(range(8))
    none_mask, _, _ = get_ltor_masks_and_position_ids(data, eod, False, False, False, build_attention_mask=False)
    assert none_mask is None


def test_glu_activations_second_half_is_the_gate():
    from epfl_megatron_b200.models.glu_activations import GLU_ACTIVATIONS
    torch.manual_seed(0)
    x = torch.randn(3, 5, 16, requires_grad=True)
    a, g = x.detach().chunk(2, dim=-1)
    want = {"liglu": a * g, "geglu": a * torch.nn.functional.gelu(g), "reglu": a * torch.relu(g),
            "swiglu": a * torch.nn.functional.silu(g)}
    for name, fn in GLU_ACTIVATIONS.items():
        y = fn(x)
        assert y.shape == (3, 5, 8)
        torch.testing.assert_close(y.detach(), want[name], rtol=1e-5, atol=1e-6)
    GLU_ACTIVATIONS["swiglu"](x).sum().backward()
    ref = x.detach().clone().requires_grad_()
    ra, rg = ref.chunk(2, dim=-1)
    (ra * torch.nn.functional.silu(rg)).sum().backward()
    torch.testing.assert_close(x.grad, ref.grad, rtol=1e-5, atol=1e-6)


def test_rope_matches_complex_rotation_of_adjacent_pairs():
    from epfl_megatron_b200.models.positional_embeddings import apply_rotary_emb, precompute_freqs_cis
    torch.manual_seed(1)
    s, b, h, d = 6, 2, 3, 8
    q, k = torch.randn(s, b, h, d), torch.randn(s, b, h, d)
    table = precompute_freqs_cis(d, 16, theta=10000.0, scaling_factor=1.0)
    pos = torch.tensor([[0, 1, 2, 0, 1, 2], [3, 4, 5, 6, 7, 8]])
    rq, rk = apply_rotary_emb(q, k, table, position_ids=pos)
    inv = 1.0 / (10000.0 ** (torch.arange(0, d, 2).float() / d))
    for bi in range(b):
        ang = torch.outer(pos[bi].float(), inv)                       # [s, d/2]
        rot = torch.polar(torch.ones_like(ang), ang)[:, None, :]      # [s, 1, d/2]
        want = torch.view_as_real(torch.view_as_complex(q[:, bi].reshape(s, h, d // 2, 2)) * rot).reshape(s, h, d)
        torch.testing.assert_close(rq[:, bi], want, rtol=1e-5, atol=1e-5)
    # linear position scaling divides the angles
    scaled = precompute_freqs_cis(d, 32, theta=10000.0, scaling_factor=2.0)
    sq, _ = apply_rotary_emb(q, k, scaled, position_ids=pos * 2)
    torch.testing.assert_close(sq, rq, rtol=1e-5, atol=1e-5)
    assert rk.shape == k.shape


def test_timers_levels_and_elapsed():
    import time as _time

    from epfl_megatron_b200.timers import Timers
    t = Timers(1, "minmax")
    t("step", log_level=0).start()
    _time.sleep(0.02)
    t("step").stop()
    deep = t("too-detailed", log_level=2)        # above the configured level: a no-op stand-in
    deep.start()
    deep.stop()
    with pytest.raises(Exception):
        deep.elapsed()
    assert 0.01 < t("step").elapsed(reset=True) < 1.0
    assert t("step").elapsed() == 0.0
    with pytest.raises(AssertionError):
        t("step", log_level=1)                   # level of an existing timer cannot change


def test_vocab_padding_and_null_tokenizer():
    from epfl_megatron_b200.tokenizer.tokenizer import _vocab_size_with_padding, build_tokenizer
    args = types.SimpleNamespace(rank=1, make_vocab_size_divisible_by=128, tensor_model_parallel_size=4)
    assert _vocab_size_with_padding(32000, args) == 32256 and _vocab_size_with_padding(32256, args) == 32256
    args = types.SimpleNamespace(rank=1, tokenizer_type="NullTokenizer", vocab_size=1000, padded_vocab_size=None,
                                 make_vocab_size_divisible_by=64, tensor_model_parallel_size=2)
    tok = build_tokenizer(args)
    assert args.padded_vocab_size % 128 == 0 and args.padded_vocab_size >= tok.vocab_size
    assert tok.detokenize(tok.tokenize("1 2 3")) == "1 2 3"


def test_sentencepiece_tokenizer_special_tokens(tmp_path):
    spm = pytest.importorskip("sentencepiece")
    from epfl_megatron_b200.tokenizer.tokenizer import _SentencePieceTokenizer
    corpus = tmp_path / "c.txt"
    corpus.write_text("\n".join(["the quick brown fox jumps over the lazy dog", "pack my box with five dozen jugs",
                                 "sphinx of black quartz judge my vow"] * 30))
    prefix = str(tmp_path / "m")
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=prefix, vocab_size=60, model_type="bpe",
                                   bos_id=1, eos_id=2, unk_id=0, pad_id=-1)
    plain = _SentencePieceTokenizer(prefix + ".model", new_tokens=False)
    tok = _SentencePieceTokenizer(prefix + ".model", vocab_extra_ids_list="<|im_start|>,<|im_end|>", new_tokens=True)
    assert tok.vocab_size > plain.vocab_size
    for name in ("<CLS>", "<SEP>", "<EOD>", "<MASK>", "<|im_start|>", "<|im_end|>"):
        assert name in tok.vocab and tok.inv_vocab[tok.vocab[name]] == name
    ids = tok.tokenize("<|im_start|>the quick fox<|im_end|>")
    assert ids[0] == tok.vocab["<|im_start|>"] and ids[-1] == tok.vocab["<|im_end|>"]
    assert all(i < plain.vocab_size for i in ids[1:-1])
    assert "the quick fox" in tok.detokenize(ids)
    assert tok.eod == tok.vocab["<EOD>"] or tok.eod == tok.eos


def test_checkpoint_file_names():
    from epfl_megatron_b200.checkpointing import get_checkpoint_names, get_checkpoint_tracker_filename
    from epfl_megatron_b200.parallel import state as mpu
    kw = dict(pipeline_parallel=False, tensor_rank=3, pipeline_rank=0)
    m, o = get_checkpoint_names("/ck", 1200, False, **kw)
    assert m == o == "/ck/iter_0001200/mp_rank_03/model_optim_rng.pt"
    m, o = get_checkpoint_names("/ck", 7, False, pipeline_parallel=True, tensor_rank=1, pipeline_rank=2)
    assert m == "/ck/iter_0000007/mp_rank_01_002/model_optim_rng.pt"
    m, _ = get_checkpoint_names("/ck", 0, False, release=True, **kw)
    assert m == "/ck/release/mp_rank_03/model_optim_rng.pt"
    mpu.set_data_parallel_rank(5) if hasattr(mpu, "set_data_parallel_rank") else None
    try:
        m, o = get_checkpoint_names("/ck", 9, True, **kw)
        assert m == "/ck/iter_0000009/mp_rank_03/model_rng.pt" and o.startswith("/ck/iter_0000009/mp_rank_03_")
        assert o.endswith("/optim.pt")
    finally:
        mpu.set_data_parallel_rank(None) if hasattr(mpu, "set_data_parallel_rank") else None
    assert get_checkpoint_tracker_filename("/ck") == "/ck/latest_checkpointed_iteration.txt"
