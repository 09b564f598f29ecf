"""Cross-flag validation of the argument namespace (megatron/arguments.py:52-345 rules, plus the places
where this framework deliberately differs), the eager softmax path, and the auxiliary subsystems
(metric sinks, signal handler, legacy buffers, fault-injection spec)."""
import json
import os
import signal

import pytest
import torch

import finetune
from epfl_megatron_b200.arguments import parse_args, validate_args

BASE = ["--num_layers", "8", "--hidden_size", "256", "--num_attention_heads", "8", "--seq_length", "128",
        "--max_position_embeddings", "128", "--micro_batch_size", "2", "--train_iters", "10", "--lr", "1e-4",
        "--tokenizer_type", "NullTokenizer"]


def _args(extra, world=8):
    a = parse_args(finetune.extra_args, args_list=BASE + extra)
    a.world_size, a.rank = world, 1       # rank 1: no banner
    return validate_args(a, {})


def test_derived_sizes_and_dtypes():
    a = _args(["--tensor_model_parallel_size", "2", "--pipeline_model_parallel_size", "2", "--bf16"])
    assert a.data_parallel_size == 2 and a.global_batch_size == 4          # mbs * dp by default
    assert a.params_dtype == torch.bfloat16 and a.accumulate_allreduce_grads_in_fp32
    assert a.ffn_hidden_size == 1024 and a.kv_channels == 32 and a.num_attention_heads_kv == 8
    assert a.encoder_seq_length == 128 and a.virtual_pipeline_model_parallel_size is None
    assert _args(["--fp16"]).params_dtype == torch.half
    with pytest.raises(AssertionError):
        _args(["--tensor_model_parallel_size", "3"])                       # 8 % 3
    with pytest.raises(AssertionError):
        _args(["--fp16", "--bf16"])


def test_sequence_parallel_rules():
    assert _args(["--sequence_parallel"]).sequence_parallel is False        # forced off at tp 1
    a = _args(["--sequence_parallel", "--tensor_model_parallel_size", "2"])
    assert a.sequence_parallel and a.async_tensor_model_parallel_allreduce is False


def test_interleaving_allowed_for_two_stages():
    a = _args(["--pipeline_model_parallel_size", "2", "--num_layers_per_virtual_pipeline_stage", "2"])
    assert a.virtual_pipeline_model_parallel_size == 2                      # 8 layers / 2 stages / 2
    with pytest.raises(AssertionError):
        _args(["--num_layers_per_virtual_pipeline_stage", "2"])             # pp == 1
    with pytest.raises(AssertionError):
        _args(["--pipeline_model_parallel_size", "2", "--num_layers_per_virtual_pipeline_stage", "3"])


def test_gqa_recompute_and_misc_rules():
    with pytest.raises(AssertionError):
        _args(["--num_attention_heads_kv", "2", "--tensor_model_parallel_size", "4"])
    with pytest.raises(AssertionError):                                     # multi-query needs tp 1
        _args(["--num_attention_heads_kv", "1", "--tensor_model_parallel_size", "4"])
    assert _args(["--num_attention_heads_kv", "1"]).num_attention_heads_kv == 1
    with pytest.raises(AssertionError):
        _args(["--recompute_granularity", "selective", "--recompute_method", "uniform"])
    with pytest.raises(AssertionError):
        _args(["--parallel_layernorm"])                                     # needs --parallel_attn
    with pytest.raises(AssertionError):
        _args(["--fp8_hybrid"])                                             # FP8 linears need --bf16
    assert _args(["--fp8_hybrid", "--bf16"]).fp8_hybrid                     # local layers run the FP8 recipe
    with pytest.raises(AssertionError):
        _args(["--fp8_hybrid", "--fp8_e4m3", "--bf16"])                     # one format
    with pytest.raises(AssertionError):
        _args(["--save", "/tmp/x"])                                         # needs --save_interval
    with pytest.raises(AssertionError):
        _args(["--min_lr", "1.0"])                                          # above --lr
    with pytest.raises(AssertionError):
        _args(["--train_samples", "100"])                                   # mixed with --train_iters
    a = _args(["--use_distributed_optimizer"])
    assert a.use_contiguous_buffers_in_local_ddp and a.DDP_impl == "local"


def test_eager_scale_mask_softmax_matches_kernel_contract():
    from epfl_megatron_b200.models.enums import AttnMaskType
    from epfl_megatron_b200.models.fused_softmax import FusedScaleMaskSoftmax
    from epfl_megatron_b200.models.utils import attention_mask_func
    torch.manual_seed(0)
    x = torch.randn(2, 3, 8, 8)
    sm = FusedScaleMaskSoftmax(False, False, AttnMaskType.padding, True, attention_mask_func, True, 0.5)
    mask = torch.zeros(2, 1, 8, 8, dtype=torch.bool)
    mask[:, :, :, 5:] = True
    y = sm(x, mask)
    assert torch.allclose(y.sum(-1), torch.ones(2, 3, 8)) and float(y[..., 5:].max()) < 1e-3
    want = torch.softmax((x * 0.5).masked_fill(mask, -10000.0), dim=-1)
    torch.testing.assert_close(y, want)
    causal = torch.triu(torch.ones(8, 8, dtype=torch.bool), 1).view(1, 1, 8, 8)
    yc = FusedScaleMaskSoftmax(False, False, AttnMaskType.causal, True, attention_mask_func, True, None)(x, causal)
    assert float(yc[0, 0, 0, 0]) == pytest.approx(1.0) and float(yc[0, 0, 2, 3:].max()) < 1e-3
    with pytest.raises(AssertionError):
        FusedScaleMaskSoftmax(True, True, AttnMaskType.causal, True, attention_mask_func, True, None)


def test_jsonl_writer_and_wandb_shim(tmp_path):
    from epfl_megatron_b200.utils.jsonl_writer import JsonlScalarWriter
    from epfl_megatron_b200.wandb_logger import WandBConfig, WandbTBShim
    w = JsonlScalarWriter(str(tmp_path / "m.jsonl"))
    w.add_scalar("lm loss", torch.tensor(2.5), 3)
    w.add_scalar("not a number", "abc", 3)
    w.add_text("args", "hello", 0)
    w.close()
    rows = [json.loads(l) for l in open(tmp_path / "m.jsonl")]
    assert rows[0]["name"] == "lm loss" and rows[0]["value"] == 2.5 and rows[0]["step"] == 3
    assert rows[1]["text"] == "hello" and len(rows) == 2
    shim = WandbTBShim(WandBConfig(project="p"))           # no key, no wandb: buffers locally
    shim.add_scalar("loss", 1.0, 1)
    shim.add_scalar("lr", 0.1, 1)
    shim.add_scalar("loss vs samples", 1.0, 1)             # the "vs samples" twins are dropped
    shim.add_scalar("loss", 0.9, 2)                        # a new step flushes the previous one
    shim.flush_all()
    assert shim.history == [(1, {"loss": 1.0, "lr": 0.1}), (2, {"loss": 0.9})]


def test_signal_handler_single_process():
    from epfl_megatron_b200.dist_signal_handler import DistributedSignalHandler
    before = signal.getsignal(signal.SIGTERM)
    with DistributedSignalHandler() as h:
        assert h.signals_received() == [False]
        os.kill(os.getpid(), signal.SIGTERM)
        assert h.signals_received() == [True]
    assert signal.getsignal(signal.SIGTERM) == before and h.release() is False


def test_legacy_memory_buffers():
    from epfl_megatron_b200.memory import RingMemBuffer, allocate_mem_buff, get_mem_buff
    buf = allocate_mem_buff("act", 16, torch.float32, track_usage=True)
    assert get_mem_buff("act") is buf
    v = buf.add(torch.arange(6, dtype=torch.float32).view(2, 3))
    assert v.shape == (2, 3) and buf.is_in_use() and buf.numel_in_use() == 6
    assert buf.get_data().tolist() == [0, 1, 2, 3, 4, 5]
    with pytest.raises(AssertionError):
        buf.add(torch.zeros(11))
    buf.reset()
    assert not buf.is_in_use()
    ring = RingMemBuffer("ring", 2, 8, torch.float32, False)
    a, b = ring.get_next_buffer(), ring.get_next_buffer()
    assert a is not b and ring.get_next_buffer() is a


def test_fault_spec():
    from epfl_megatron_b200.utils.fault import maybe_inject, parse
    assert parse("nan:3:17") == ("nan", 3, 17) and parse(None) is None and parse("") is None
    lin = torch.nn.Linear(2, 2)
    lin.weight.main_grad = torch.zeros(2, 2)
    maybe_inject("nan:0:5", 4, [lin])
    assert torch.isfinite(lin.weight.main_grad).all()
    maybe_inject("nan:0:5", 5, [lin])
    assert torch.isnan(lin.weight.main_grad.view(-1)[0])
