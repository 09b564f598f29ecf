"""RMSNorm backward without a scratch argument (``b200_rmsnorm_bwd_noalloc``).

Some users of the kernel library cannot obtain device memory at the moment they need the
RMSNorm backward: a launch sequence that is being recorded must not call the allocator, a
host application may have handed the GPU's memory out already, and a plug-in that is only
given input and output pointers has no allocator at all. For them the library carries its
own fixed 148 x 8192 fp32 scratch array on the device. This test drives that entry point
through the plain C ABI (ctypes on the shared object, exactly what a C caller links
against) and requires the same dx / dw as the regular entry point, which is given a
torch-allocated scratch buffer, on two calls in a row with different shapes -- the second
call finds the first call's leftovers in the shared array and must not be disturbed by them.

The file name sorts last on purpose: a run that stops at the first failure has then
already executed every other test.
"""
import ctypes as C
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = Path(__file__).resolve().parent.parent / "epfl_megatron_b200" / "ops" / "_lib" / "libb200kernels.so"
vp, i32, f32 = C.c_void_p, C.c_int, C.c_float


def _load():
    lib = C.CDLL(str(LIB))
    lib.b200_norm_bwd_grid.argtypes, lib.b200_norm_bwd_grid.restype = [i32], i32
    lib.b200_rmsnorm_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, f32, vp]
    lib.b200_rmsnorm_fwd.restype = i32
    lib.b200_rmsnorm_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp]
    lib.b200_rmsnorm_bwd.restype = i32
    lib.b200_rmsnorm_bwd_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp]
    lib.b200_rmsnorm_bwd_noalloc.restype = i32
    lib.b200_rmsnorm_bwd_noalloc_max_h.argtypes, lib.b200_rmsnorm_bwd_noalloc_max_h.restype = [], i32
    return lib


def _close(a, b, atol, rtol, what):
    err = (a.float() - b.float()).abs()
    tol = atol + rtol * b.float().abs()
    assert bool((err <= tol).all()), f"{what}: max abs err {float(err.max())}"
    assert bool(torch.isfinite(a.float()).all()), f"{what}: non-finite values"


def _one_case(lib, M, H, seed, with_dres, accumulate):
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(seed)
    x = torch.randn(M, H, device=dev, generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(H, device=dev, generator=g)).bfloat16()
    dy = torch.randn(M, H, device=dev, generator=g).bfloat16()
    dres = torch.randn(M, H, device=dev, generator=g).bfloat16() if with_dres else None
    y = torch.empty_like(x)
    rstd = torch.empty(M, device=dev, dtype=torch.float32)
    assert lib.b200_rmsnorm_fwd(x.data_ptr(), None, w.data_ptr(), y.data_ptr(), None, rstd.data_ptr(),
                                M, H, 1e-5, stream) == 0
    dw0 = torch.randn(H, device=dev, generator=g) if accumulate else torch.zeros(H, device=dev)
    dp = dres.data_ptr() if dres is not None else None
    # existing entry point: caller-provided scratch
    scratch = torch.empty(lib.b200_norm_bwd_grid(M), H, device=dev, dtype=torch.float32)
    dx_ref, dw_ref = torch.empty_like(x), dw0.clone()
    assert lib.b200_rmsnorm_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), rstd.data_ptr(), dp,
                                dx_ref.data_ptr(), dw_ref.data_ptr(), scratch.data_ptr(), M, H,
                                int(accumulate), stream) == 0
    # new entry point: no scratch argument, nothing allocated inside
    dx, dw = torch.empty_like(x), dw0.clone()
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    # the driver's own count of free bytes sees a cudaMalloc inside the library, which torch's
    # allocator statistics would not (the one-time symbol look-up of the first call on a device
    # happened in _warm_up, before this snapshot)
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.b200_rmsnorm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), w.data_ptr(), rstd.data_ptr(), dp,
                                      dx.data_ptr(), dw.data_ptr(), M, H, int(accumulate), stream)
    assert rc == 0, f"b200_rmsnorm_bwd_noalloc returned {rc}"
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was taken during the call"
    _close(dx, dx_ref, 2e-2, 2e-2, f"dx M={M} H={H}")
    # dw: both entry points sum the same fp32 partial rows in the same order; against the fp32
    # restatement the error is rounding only, so a dropped partial row (1/grid of the sum) shows
    dw_fp32 = dw0 + (dy.float() * (x.float() * rstd[:, None]).bfloat16().float()).sum(0)
    scale = float(dw_fp32.abs().max())
    _close(dw, dw_ref, 1e-4 * scale, 1e-4, f"dw vs regular entry point M={M} H={H}")
    _close(dw, dw_fp32, 2e-3 * scale, 2e-3, f"dw vs fp32 restatement M={M} H={H}")


def _warm_up(lib):
    """First call on this device: the entry point looks up the address of its array (documented:
    make this call outside any stream capture)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.zeros(8, 256, device=dev).bfloat16()
    r = torch.ones(8, device=dev)
    o, dw = torch.empty_like(t), torch.zeros(256, device=dev)
    assert lib.b200_rmsnorm_bwd_noalloc(t.data_ptr(), t.data_ptr(), t[0].data_ptr(), r.data_ptr(), None,
                                        o.data_ptr(), dw.data_ptr(), 8, 256, 0,
                                        torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    lib = _load()
    _warm_up(lib)
    # first call: tall input, all 148 partial rows get written
    _one_case(lib, M=2048, H=4096, seed=1, with_dres=False, accumulate=False)
    # second call straight after it: fewer rows than CTAs and a wider H, so the array still
    # holds rows from the first call that this call must neither read nor need cleared
    _one_case(lib, M=96, H=8192, seed=2, with_dres=True, accumulate=True)
    _one_case(lib, M=333, H=1024, seed=3, with_dres=False, accumulate=True)


def test_noalloc_rejects_sizes_beyond_its_array():
    lib = _load()
    max_h = lib.b200_rmsnorm_bwd_noalloc_max_h()
    assert max_h == 8192
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.zeros(8, device=dev)
    rc = lib.b200_rmsnorm_bwd_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), None,
                                      t.data_ptr(), t.data_ptr(), 4, max_h + 8, 0,
                                      torch.cuda.current_stream().cuda_stream)
    assert rc == -3, "H beyond the built-in array must be refused with B200_ERR_TOO_LARGE"
    torch.cuda.synchronize()
