"""Builds the native code in-tree.

Two shared objects are produced next to this file:

* ``_lib/libb200kernels.so``  -- every CUDA kernel (sm_100a only), exported with a
  plain C ABI declared in ``include/b200_kernels.h`` and bound through ctypes.
* ``_lib/libb200data.so``     -- CPU index builders for the dataset pipeline
  (the job ``megatron/data/helpers.cpp`` does in the reference).

``python -m epfl_megatron_b200.ops.build`` rebuilds when a source is newer than the
library. nvcc cross-compiles without a GPU, so this also runs on the CPU-only box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_ROOT = Path(__file__).resolve().parent.parent
CSRC = PKG_ROOT / "csrc"
LIB_DIR = Path(__file__).resolve().parent / "_lib"
KERNEL_LIB = LIB_DIR / "libb200kernels.so"
DATA_LIB = LIB_DIR / "libb200data.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; set NVCC")
    return cand


def _newer(srcs, target: Path) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in srcs)


def cuda_sources():
    return sorted(CSRC.glob("*.cu"))


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    LIB_DIR.mkdir(exist_ok=True)
    srcs = cuda_sources()
    deps = list(srcs) + list(CSRC.glob("*.cuh"))
    if not force and not _newer(deps, KERNEL_LIB):
        return KERNEL_LIB
    obj_dir = LIB_DIR / "obj"
    obj_dir.mkdir(exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: Path) -> Path:
        obj = obj_dir / (src.stem + ".o")
        if force or _newer([src] + list(CSRC.glob("*.cuh")), obj):
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
                print(" ".join(cmd), flush=True)
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
            if verbose:
                print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    cmd = [nvcc, "-shared", "-o", str(KERNEL_LIB), *map(str, objs), "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return KERNEL_LIB


def build_data_helpers(force: bool = False) -> Path:
    LIB_DIR.mkdir(exist_ok=True)
    src = CSRC / "data_helpers.cpp"
    if not force and not _newer([src], DATA_LIB):
        return DATA_LIB
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    cmd = [cxx, "-O3", "-std=c++17", "-shared", "-fPIC", str(src), "-o", str(DATA_LIB)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed:\n{r.stdout}\n{r.stderr}")
    return DATA_LIB


def build_all(force: bool = False, verbose: bool = False):
    return build_kernels(force, verbose), build_data_helpers(force)


if __name__ == "__main__":
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv
    if "--data-only" in sys.argv:       # CPU index builders only (no nvcc needed): tokenisation hosts
        print("built", build_data_helpers(force))
    else:
        for p in build_all(force, verbose):
            print("built", p)
