"""ctypes binding of the C++ index builders (csrc/data_helpers.cpp), with NumPy fallbacks
so a missing compiler never blocks a run. Same entry points as the reference's pybind
module ``megatron.data.helpers`` (helpers.cpp:696-701)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

_LIB = None
_TRIED = False
LIB_PATH = Path(__file__).resolve().parent.parent / "ops" / "_lib" / "libb200data.so"


def _lib():
    global _LIB, _TRIED
    if _LIB is not None or _TRIED:
        return _LIB
    _TRIED = True
    try:
        if not LIB_PATH.exists():
            from ..ops.build import build_data_helpers
            build_data_helpers()
        lib = C.CDLL(str(LIB_PATH))
        lib.b200_build_sample_idx.restype = C.c_int64
        lib.b200_build_sample_idx.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p]
        lib.b200_build_blending_indices.restype = None
        lib.b200_build_blending_indices.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64]
        lib.b200_build_mapping.restype = C.c_int64
        lib.b200_build_mapping.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int64, C.c_int32,
                                           C.c_double, C.c_int32, C.c_int32, C.c_void_p]
        lib.b200_build_blocks_mapping.restype = C.c_int64
        lib.b200_build_blocks_mapping.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32,
                                                  C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
        _LIB = lib
    except Exception as e:  # noqa: BLE001
        print(f"WARNING: native data helpers unavailable ({e}); using the NumPy fallbacks", flush=True)
    return _LIB


def native_available():
    return _lib() is not None


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def build_sample_idx_py(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch):
    """Reference algorithm in Python (megatron/data/gpt_dataset.py:445-491 behaviour)."""
    num_samples = (num_epochs * tokens_per_epoch - 1) // seq_length
    out = np.zeros([num_samples + 1, 2], dtype=np.int32)
    di, off = 0, 0
    for row in range(1, num_samples + 1):
        need = seq_length + 1
        while need > 0:
            avail = int(sizes[doc_idx[di]]) - off
            if avail >= need:
                off += need - 1
                need = 0
            else:
                need -= avail
                di += 1
                off = 0
        out[row] = (di, off)
    return out


def build_sample_idx(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch):
    num_samples = (num_epochs * tokens_per_epoch - 1) // seq_length
    lib = _lib()
    if lib is None:
        return build_sample_idx_py(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    doc_idx = np.ascontiguousarray(doc_idx, dtype=np.int32)
    out = np.zeros([num_samples + 1, 2], dtype=np.int32)
    rows = lib.b200_build_sample_idx(_p(sizes), _p(doc_idx), len(doc_idx), int(seq_length), int(num_samples), _p(out))
    assert rows == num_samples + 1, "ran out of tokens while building the sample index"
    return out


def build_blending_indices(dataset_index, dataset_sample_index, weights, num_datasets, size, verbose=False):
    weights = np.ascontiguousarray(weights, dtype=np.float64)
    lib = _lib()
    if lib is None:
        taken = np.zeros(num_datasets, dtype=np.int64)
        for i in range(size):
            err = weights * max(i, 1) - taken
            d = int(np.argmax(err))
            dataset_index[i] = d
            dataset_sample_index[i] = taken[d]
            taken[d] += 1
        return
    assert dataset_index.dtype == np.uint8 and dataset_sample_index.dtype == np.int64
    lib.b200_build_blending_indices(_p(dataset_index), _p(dataset_sample_index), _p(weights), int(num_datasets),
                                    int(size))


def build_mapping(docs, sizes, num_epochs, max_num_samples, max_seq_length, short_seq_prob, seed, verbose,
                  min_num_sent):
    lib = _lib()
    assert lib is not None, "build_mapping needs the native helpers (python -m epfl_megatron_b200.ops.build)"
    docs = np.ascontiguousarray(docs, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    args = (_p(docs), len(docs) - 1, _p(sizes), int(num_epochs), int(max_num_samples), int(max_seq_length),
            float(short_seq_prob), int(seed), int(min_num_sent))
    n = lib.b200_build_mapping(*args, None)
    out = np.zeros([n, 3], dtype=np.int64)
    lib.b200_build_mapping(*args, _p(out))
    return out


def build_blocks_mapping(docs, sizes, title_sizes, num_epochs, max_num_samples, max_seq_length, seed, verbose,
                         use_one_sent_blocks):
    lib = _lib()
    assert lib is not None, "build_blocks_mapping needs the native helpers"
    docs = np.ascontiguousarray(docs, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    title_sizes = np.ascontiguousarray(title_sizes, dtype=np.int32)
    args = (_p(docs), len(docs) - 1, _p(sizes), _p(title_sizes), int(num_epochs), int(max_num_samples),
            int(max_seq_length), int(seed), int(bool(use_one_sent_blocks)))
    n = lib.b200_build_blocks_mapping(*args, None)
    out = np.zeros([n, 4], dtype=np.int64)
    lib.b200_build_blocks_mapping(*args, _p(out))
    return out
