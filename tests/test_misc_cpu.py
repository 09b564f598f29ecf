"""Small CPU-only pieces: the vestigial vision data utilities and the unpickling stubs."""
import io
import pickle
import random

import pytest


def _write_images(root, classes=("ant", "bee"), per_class=3):
    from PIL import Image
    for ci, c in enumerate(classes):
        d = root / c
        d.mkdir()
        for i in range(per_class):
            Image.new("RGB", (24, 16), (40 * ci + 10, 20 * i, 200)).save(d / f"{i}.png")
        (d / "notes.txt").write_text("not an image")


def test_image_folder(tmp_path):
    pytest.importorskip("PIL")
    from epfl_megatron_b200.data.image_folder import ImageFolder, is_image_file
    _write_images(tmp_path)
    ds = ImageFolder(str(tmp_path))
    assert ds.classes == ["ant", "bee"] and len(ds) == 6
    img, target = ds[4]
    assert img.size == (24, 16) and target == 1
    assert is_image_file("x.JPG") and not is_image_file("x.txt")
    half = ImageFolder(str(tmp_path), classes_fraction=0.5)
    assert half.classes == ["ant"] and len(half) == 3
    few = ImageFolder(str(tmp_path), data_per_class_fraction=0.34, transform=lambda im: im.size)
    assert len(few) == 2 and few[0][0] == (24, 16)


def test_autoaugment_policy():
    pytest.importorskip("PIL")
    from PIL import Image
    from epfl_megatron_b200.data.autoaugment import ImageNetPolicy, SubPolicy
    random.seed(3)
    img = Image.new("RGB", (32, 32), (120, 30, 200))
    pol = ImageNetPolicy()
    assert len(pol.policies) == 25
    for _ in range(60):
        out = pol(img)
        assert out.size == img.size and out.mode == "RGB"
    # every operation runs at full probability
    for op in ("shearX", "shearY", "translateX", "translateY", "rotate", "color", "posterize", "solarize",
               "contrast", "sharpness", "brightness", "autocontrast", "equalize", "invert"):
        assert SubPolicy(op, 1.0, 9, "invert", 0.0, 0)(img).size == img.size
    with pytest.raises(ValueError):
        SubPolicy("blur", 0.5, 1, "invert", 0.5, 1)


def test_deprecated_loss_scaler_unpickles_under_reference_path():
    import megatron  # noqa: F401  (installs the megatron.* aliases)
    from megatron.fp16_deprecated.loss_scaler import DynamicLossScaler, LossScaler
    assert DynamicLossScaler.__module__.endswith("fp16_deprecated.loss_scaler")
    blob = pickle.dumps({"a": LossScaler(4.0), "b": DynamicLossScaler(init_scale=8)})
    back = pickle.load(io.BytesIO(blob))
    assert back["a"].cur_scale == 4.0 and back["b"].cur_scale == 8


def test_distrib_optimizer_module_path():
    from megatron.optimizer.distrib_optimizer import DistributedOptimizer, Range
    r = Range(8, 24).normalize(100)
    assert (r.start, r.end, r.size) == (100, 116, 16) and DistributedOptimizer.__name__ == "DistributedOptimizer"


def test_fused_dp_buffer_order_rotates_bucket_owners():
    """Bucket-sized runs of parameters (in gradient production order) are dealt round-robin over the
    dp regions of the buffer, each run starting a bucket (models/distributed.py::_buffer_order)."""
    import types

    import torch
    from epfl_megatron_b200.models.distributed import DistributedDataParallel as DDP

    params = [torch.nn.Parameter(torch.zeros(n)) for n in (100, 30, 80, 60, 10, 120, 40, 90, 70)]
    produced = list(enumerate(params))
    fake = types.SimpleNamespace(bucket_bytes=100 * 4)
    flat = DDP._buffer_order(fake, produced, lambda p: torch.float32, 1)
    assert [i for i, _, _ in flat] == list(range(9)) and not any(c for _, _, c in flat)
    out = DDP._buffer_order(fake, produced, lambda p: torch.float32, 2)
    # runs by >= 400 bytes: [0] [1,2] [3,4,5] [6,7] [8]; lane 0 takes runs 0,2,4, lane 1 runs 1,3
    assert [i for i, _, _ in out] == [0, 3, 4, 5, 8, 1, 2, 6, 7]
    assert [i for i, _, c in out if c] == [0, 3, 8, 1, 6]
    assert sorted(i for i, _, _ in out) == list(range(9))


def test_bench_argument_sets_of_all_baseline_configs_validate(monkeypatch):
    """bench.py's flag sets for the five BASELINE.json configurations parse and pass validate_args."""
    import types

    import bench
    import finetune
    from epfl_megatron_b200.arguments import parse_args, validate_args

    def ns(**k):
        d = dict(model="llama2-7b", steps=2, warmup=3, seq=4096, micro_batch=1, num_microbatches=4, recompute=False,
                 debug_layers=0, tp=0, pp=1, vpp_layers=0)
        d.update(k)
        return types.SimpleNamespace(**d)

    monkeypatch.setenv("WORLD_SIZE", "8")
    monkeypatch.setenv("RANK", "0")
    cases = [(ns(), (2, 4), {}),
             (ns(model="llama2-13b", tp=4, pp=2, vpp_layers=5), (4, 1), {"virtual_pipeline_model_parallel_size": 4}),
             (ns(model="falcon-40b", tp=8, seq=2048), (8, 1), {"parallel_attn": True, "use_rms_norm": False}),
             (ns(model="llama2-70b", tp=4, pp=2, recompute=True), (4, 1), {"recompute_granularity": "full"})]
    for a, layout, expect in cases:
        tp, dp = bench.parallel_layout(8, a.model, a.tp, a.pp)
        assert (tp, dp) == layout
        argv, gbs = bench.build_arg_list(a, tp, dp)
        assert gbs == 4 * dp
        args = parse_args(finetune.extra_args, args_list=argv)
        args.world_size, args.rank = 8, 0
        args = validate_args(args, {})
        assert args.sequence_parallel and args.use_distributed_optimizer and args.bf16
        for k, v in expect.items():
            assert getattr(args, k) == v, (a.model, k, getattr(args, k))


def test_example_scripts_only_use_flags_the_parsers_know():
    """Every --flag in examples/**/*.sh is an option of the training parser, of one of the tool / task
    command lines, of the launcher, or a short flag the wrapper script defines itself."""
    import glob
    import os
    import re

    import finetune
    from epfl_megatron_b200.arguments import build_base_parser
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    parser = finetune.extra_args(build_base_parser())
    known = {o for a in parser._actions for o in a.option_strings}
    for f in ["tasks/main.py", "tasks/msdp/main.py", "tasks/msdp/preprocessing.py", "tools/run_text_generation_server.py",
              "tools/checkpoint_util.py", "tools/checkpoint_loader_megatron.py", "tools/checkpoint_saver_megatron.py",
              "tools/preprocess_data.py", "tools/merge_datasets.py", "tools/push_to_hub.py", "bench.py",
              "weights2megatron/weights2megatron.py", "weights2megatron/megatron2hf.py", "verify_correctness.py",
              "pretrain_ict.py", "pretrain_t5.py", "pretrain_bert.py"]:
        src = open(os.path.join(root, f)).read()
        for m in re.finditer(r'add_argument\(\s*"(--[A-Za-z0-9_-]+)"(?:,\s*"(--[A-Za-z0-9_-]+)")?', src):
            known.update(x for x in m.groups() if x)
    launcher = {"--nproc_per_node", "--nproc-per-node", "--nnodes", "--node_rank", "--node-rank", "--master_addr",
                "--master_port", "--master-addr", "--master-port", "--nodes", "--ntasks-per-node", "--exclusive",
                "--output", "--standalone", "--local-addr", "--help"}
    scripts = sorted(glob.glob(os.path.join(root, "examples", "**", "*.sh"), recursive=True))
    assert len(scripts) >= 35
    unknown = {}
    for f in scripts:
        txt = open(f).read()
        own = set(re.findall(r"(--[a-z-]+)\)", txt))          # case-statement flags of wrapper scripts
        for tok in set(re.findall(r"(?<![\w-])(--[A-Za-z][A-Za-z0-9_-]*)", txt)):
            if tok not in known and tok not in launcher and tok not in own:
                unknown.setdefault(os.path.relpath(f, root), set()).add(tok)
    assert not unknown, unknown


def test_fused_kernels_surface_matches_torch():
    """``megatron.fused_kernels``: the reference's five extension names with their call signatures
    (megatron/fused_kernels/tests/test_fused_kernels.py style checks, CPU restatement path)."""
    import torch
    from megatron import fused_kernels as fk
    fk.load(None)
    torch.manual_seed(0)
    x = torch.randn(2, 3, 8, 16)
    mask = torch.rand(2, 1, 8, 16) < 0.3
    mask[0, 0, 2] = True                      # a fully masked row must come out as zeros, not NaN
    y = fk.scaled_masked_softmax_cuda.forward(x, mask, 0.5)
    want = torch.softmax((x * 0.5).masked_fill(mask, -10000.0), -1)
    want[0, :, 2] = 0
    assert torch.allclose(y, want, atol=1e-5)
    g = torch.randn_like(y)
    xr = x.clone().requires_grad_()
    yr = torch.softmax((xr * 0.5).masked_fill(mask, float("-inf")).nan_to_num(neginf=-1e30), -1)
    dx = fk.scaled_masked_softmax_cuda.backward(g, y, 0.5)
    gy = g * y
    assert torch.allclose(dx, 0.5 * (gy - y * gy.sum(-1, keepdim=True)), atol=1e-5)
    tri = fk.scaled_upper_triang_masked_softmax_cuda.forward(torch.randn(6, 8, 8), 1.0)
    assert torch.allclose(tri.sum(-1), torch.ones(6, 8), atol=1e-5) and float(tri.triu(1).abs().max()) == 0.0
    assert torch.allclose(fk.scaled_softmax_cuda.forward(x, 2.0), torch.softmax(2.0 * x, -1), atol=1e-5)
    # reference sizing helper: 128 threads per block, 2 rows per warp up to 128 keys
    assert fk.scaled_masked_softmax_cuda.get_batch_per_block(8, 16, 2, 3) == (128 // 16) * 2
    assert fk.scaled_masked_softmax_cuda.get_batch_per_block(128, 2048, 2, 16) == 4
    # LayerNorm forward / backward against autograd
    h = torch.randn(5, 32, requires_grad=True)
    w, b = torch.randn(32, requires_grad=True), torch.randn(32, requires_grad=True)
    out, mean, invvar = fk.fused_mix_prec_layer_norm_cuda.forward_affine(h.detach(), (32,), w.detach(), b.detach(), 1e-5)
    ref_out = torch.nn.functional.layer_norm(h, (32,), w, b, 1e-5)
    assert torch.allclose(out, ref_out, atol=1e-5)
    go = torch.randn_like(out)
    ref_out.backward(go)
    dx, dg, db = fk.fused_mix_prec_layer_norm_cuda.backward_affine(go, mean, invvar, h.detach(), (32,), w.detach(),
                                                                   b.detach(), 1e-5)
    assert torch.allclose(dx, h.grad, atol=1e-4) and torch.allclose(dg, w.grad, atol=1e-4)
    assert torch.allclose(db, b.grad, atol=1e-4)
    # weight gradient accumulated into an fp32 buffer
    inp, gout, main = torch.randn(4, 6, 8), torch.randn(4, 6, 10), torch.ones(10, 8)
    fk.fused_dense_cuda.wgrad_gemm_accum_fp32(inp, gout, main)
    assert torch.allclose(main, 1 + gout.reshape(-1, 10).t() @ inp.reshape(-1, 8), atol=1e-4)


def test_fp8_delayed_scaling_bookkeeping():
    """ops/fp8.advance_step: scale = fp8_max / amax / 2**margin from the most recent amax or the max over the
    history; amax of zero or a non-finite scale keeps the old scale; histories roll; interval is honoured."""
    import torch
    from epfl_megatron_b200.ops import fp8
    fp8.reset()
    try:
        r = fp8.Recipe(hybrid=True, margin=1, interval=1, history_len=3, algo="max")
        pool = fp8._pool(torch.device("cpu"), r)
        fp8._N = 2
        pool["hist"][0, :, 0] = torch.tensor([2.0, 4.0, 8.0])
        pool["hist"][0, :, 1] = torch.tensor([7.0, 1.0, 1.0])
        pool["hist"][1, :, 0] = torch.tensor([0.0, float("inf"), 3.5])
        fp8.advance_step()
        want0 = torch.tensor([448.0 / 7.0, 448.0 / 4.0, 57344.0 / 8.0]) / 2.0
        assert torch.allclose(pool["scale"][0], want0)
        assert torch.allclose(pool["scale_inv"][0], 1.0 / want0)
        assert pool["scale"][1, 0] == 1.0 and pool["scale"][1, 1] == 1.0           # amax 0 / inf: unchanged
        assert torch.allclose(pool["scale"][1, 2], torch.tensor(57344.0 / 3.5 / 2.0))
        assert torch.equal(pool["hist"][0, :, 1], torch.tensor([2.0, 4.0, 8.0]))   # rolled
        assert float(pool["hist"][:2, :, 0].abs().sum()) == 0.0                    # fresh slot for the next step
        assert fp8._STEP == 1
    finally:
        fp8.reset()


def test_fused_reduce_gemm_tile_order_is_a_permutation_with_lag():
    """tp 2 reduce-scatter / all-reduce GEMM: the interleaved order (csrc/gemm_comm_tcgen05.cu) visits every tile
    once, starts with `lag` remote tiles, ends with `lag` own tiles, and reaches own tile k at least `lag`
    positions after the peer (which walks the same sequence) produced the matching remote tile k."""
    import ctypes
    from pathlib import Path
    so = Path(__file__).resolve().parent.parent / "epfl_megatron_b200" / "ops" / "_lib" / "libb200kernels.so"
    if not so.exists():
        import pytest
        pytest.skip("native library not built")
    lib = ctypes.CDLL(str(so))
    f = lib.b200_comm_interleaved_tile
    f.argtypes, f.restype = [ctypes.c_int] * 3, ctypes.c_int
    for half, pairs in ((128, 74), (64, 74), (352, 74), (8, 74), (40, 16)):
        lag = min(pairs, half)
        seq = [f(i, half, lag) for i in range(2 * half)]
        assert sorted(seq) == list(range(2 * half))
        assert all(t < half for t in seq[:lag]) and all(t >= half for t in seq[2 * half - lag:])
        pos = {t: i for i, t in enumerate(seq)}
        assert min(pos[half + k] - pos[k] for k in range(half)) >= lag
        # remote and own tiles alternate in the middle: the link is never asked for more than every other tile
        mid = seq[lag:2 * half - lag]
        assert all((a < half) != (b < half) for a, b in zip(mid, mid[1:]))


def test_c_abi_header_sources_and_bindings_agree():
    """Every exported entry point (B200_API in csrc/*.cu) is declared in include/b200_kernels.h, and every
    ctypes binding of ops/lib.py names an entry point that exists."""
    import re
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    src = set()
    for f in (root / "epfl_megatron_b200" / "csrc").glob("*.cu"):
        src |= set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", f.read_text()))
    hdr = set(re.findall(r"\b(b200_\w+)\s*\(", (root / "include" / "b200_kernels.h").read_text()))
    bound = set(re.findall(r'"(b200_\w+)"\s*:', (root / "epfl_megatron_b200" / "ops" / "lib.py").read_text()))
    assert len(src) > 60
    assert not (src - hdr), f"exported but not declared in the header: {sorted(src - hdr)}"
    assert not (hdr - src), f"declared but not defined: {sorted(hdr - src)}"
    assert not (bound - src), f"bound in lib.py but not defined: {sorted(bound - src)}"


def test_fp8_gpu_path_python_side_with_recording_library(monkeypatch):
    """The non-emulated FP8 path (the one a GPU runs) with the native calls recorded instead of executed:
    every Python line between the autograd function and the C entry points runs, and the arguments are the
    ones the entry points document (operand roles, formats, strides, output modes, scale pointers)."""
    import torch
    from epfl_megatron_b200.ops import fp8
    fp8.reset()
    calls = []
    monkeypatch.setattr(fp8, "_EMULATE", False)
    monkeypatch.setattr(fp8.L, "call", lambda name, *a, **k: calls.append((name, a)) or 0)
    monkeypatch.setattr(fp8.L, "stream_ptr", lambda: 0)
    try:
        recipe = fp8.Recipe(hybrid=True, history_len=2, algo="max")
        w = torch.zeros(48, 32, dtype=torch.bfloat16)
        w._b200_fp8 = meta = fp8.Meta(torch.device("cpu"), recipe)
        x = torch.zeros(64, 32, dtype=torch.bfloat16)
        dy = torch.zeros(64, 48, dtype=torch.bfloat16)
        y = fp8.forward_gemm(x, w)
        assert y.shape == (64, 48) and y.dtype == torch.bfloat16
        q_x, q_w, g_fwd = calls
        assert q_x[0] == q_w[0] == "b200_fp8_quantize" and g_fwd[0] == "b200_gemm_fp8"
        assert q_x[1][3:6] == (64, 32, 32) and q_x[1][2] is None and q_x[1][8] == 0          # no transpose, E4M3
        assert q_w[1][3:6] == (48, 32, 32) and q_w[1][2] is not None                          # weight + transpose
        assert q_x[1][6] == meta.scale_ptr(fp8.X) and q_x[1][7] == meta.amax_ptr(fp8.X)
        assert g_fwd[1][3:8] == (64, 48, 32, 0, 0) and g_fwd[1][8:12] == (32, 32, 48, 0)
        assert g_fwd[1][12:14] == (meta.inv_ptr(fp8.X), meta.inv_ptr(fp8.W))
        calls.clear()
        dx, dyT8 = fp8.dgrad_gemm(dy, w)                       # weights cached: one cast + one GEMM
        assert [c[0] for c in calls] == ["b200_fp8_quantize", "b200_gemm_fp8"]
        assert dx.shape == (64, 32) and dyT8.shape == (48, 64)
        assert calls[0][1][8] == 1 and calls[0][1][6] == meta.scale_ptr(fp8.G)                # E5M2 gradients
        assert calls[1][1][3:8] == (64, 32, 48, 1, 0) and calls[1][1][8:12] == (48, 48, 32, 0)
        assert calls[1][1][12:14] == (meta.inv_ptr(fp8.G), meta.inv_ptr(fp8.W))
        calls.clear()
        main = torch.zeros(48, 32)
        fp8.wgrad_gemm(dyT8, x, w, main, accumulate=True)
        assert calls[0][1][7] is None, "the re-cast of x for the weight gradient must not touch the amax history"
        assert calls[1][1][3:8] == (48, 32, 64, 1, 0) and calls[1][1][11] == 2
        assert calls[1][1][12:14] == (meta.inv_ptr(fp8.G), meta.inv_ptr(fp8.X))
        calls.clear()
        fp8.advance_step()
        fp8.forward_gemm(x, w)                                 # new optimizer step: the weight is cast again
        assert [c[0] for c in calls].count("b200_fp8_quantize") == 2
    finally:
        fp8.reset()
