"""ctypes binding of libspeq_cuda.so (include/speq_cuda.h) -- the same C ABI the C++ `speq` host binary links.

Nothing here computes: every call lands in the CUDA library.  If the library has not been built the import of the
handle raises; there is no Python / CPU fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
LIB_PATH = ROOT / "lib" / "libspeq_cuda.so"

OK, EINVAL, ECUDA, EIO, ENOMEM, ESTATE = 0, -1, -2, -3, -4, -5
BUILD_KEEP_SA = 1
BUILD_DEVICE_SA = 2
ENC_ASCII, ENC_RANK = 0, 1
MODE_GLOBAL, MODE_LOCAL = 0, 1
ALGO_ROLLING, ALGO_DIRECT, ALGO_ROLLING64, ALGO_KTABLE, ALGO_AUTO, ALGO_KPAIR = 0, 1, 2, 3, 4, 5

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
i32p = C.POINTER(C.c_int32)
f64p = C.POINTER(C.c_double)


class IndexInfo(C.Structure):
    _fields_ = [("n", C.c_uint64), ("n_texts", C.c_uint32), ("n_groups", C.c_uint32), ("C", C.c_uint64 * 8),
                ("n_exceptions", C.c_uint64), ("device_bytes", C.c_uint64)]


class Batch(C.Structure):
    _fields_ = [("bases", C.c_void_p), ("inv", C.c_void_p), ("qual", C.c_void_p), ("grp_off", C.c_void_p),
                ("len", C.c_void_p), ("n_reads", C.c_uint64), ("n_groups32", C.c_uint64)]


class Params(C.Structure):
    _fields_ = [("k", C.c_uint32), ("phred_cutoff", C.c_uint32), ("mode", C.c_int32), ("paired", C.c_int32),
                ("record_multiplicity", C.c_int32), ("algorithm", C.c_uint32), ("stream", C.c_void_p)]


class Counts(C.Structure):
    _fields_ = [("total_kmers", C.c_uint64), ("ambiguous_reads", C.c_uint64), ("lookups", C.c_uint64),
                ("reads", C.c_uint64), ("kernel_launches", C.c_uint64), ("steps_executed", C.c_uint64),
                ("kmers_hit", C.c_uint64), ("occurrences", C.c_uint64), ("occ_sectors", C.c_uint64)]


# every symbol include/speq_cuda.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "speqg_last_error": (C.c_char_p, []),
    "speqg_abi_version": (C.c_int, []),
    "speqg_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "speqg_index_build": (C.c_int, [C.POINTER(C.c_void_p), u64p, C.c_uint32, i32p, C.c_uint32, C.c_uint32,
                                    C.POINTER(C.c_void_p)]),
    "speqg_index_save": (C.c_int, [C.c_void_p, C.c_char_p]),
    "speqg_index_load": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "speqg_index_save_at": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64]),
    "speqg_index_load_at": (C.c_int, [C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p)]),
    "speqg_refidx_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "speqg_refidx_reference_path": (C.c_char_p, [C.c_void_p]),
    "speqg_refidx_file_time": (C.c_int64, [C.c_void_p, C.c_int]),
    "speqg_refidx_size": (C.c_uint64, [C.c_void_p]),
    "speqg_refidx_n_groups": (C.c_uint32, [C.c_void_p]),
    "speqg_refidx_group_name": (C.c_char_p, [C.c_void_p, C.c_uint32]),
    "speqg_refidx_group_count": (C.c_int32, [C.c_void_p, C.c_uint32]),
    "speqg_refidx_n_scaffolds": (C.c_uint32, [C.c_void_p]),
    "speqg_refidx_group_of_scaffold": (C.c_int32, [C.c_void_p, C.c_uint32]),
    "speqg_refidx_n_texts": (C.c_uint32, [C.c_void_p]),
    "speqg_refidx_text": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), u64p]),
    "speqg_refidx_build_index": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p)]),
    "speqg_refidx_close": (None, [C.c_void_p]),
    "speqg_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "speqg_host_free": (None, [C.c_void_p]),
    "speqg_scan_wait_copies": (C.c_int, [C.c_void_p]),
    "speqg_index_get_info": (C.c_int, [C.c_void_p, C.POINTER(IndexInfo)]),
    "speqg_index_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "speqg_index_upload": (C.c_int, [C.c_void_p, C.c_int]),
    "speqg_index_free": (None, [C.c_void_p]),
    "speqg_pack_groups": (C.c_uint64, [C.c_void_p, C.c_uint64]),
    "speqg_pack_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.c_int, C.c_int,
                                   C.POINTER(Batch)]),
    "speqg_scan_open": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Params), C.POINTER(C.c_void_p)]),
    "speqg_scan_submit_packed": (C.c_int, [C.c_void_p, C.POINTER(Batch)]),
    "speqg_scan_submit_raw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]),
    "speqg_scan_submit_fastq": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, u64p, u64p]),
    "speqg_scan_submit_fastq_pair": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, u64p, u64p]),
    "speqg_scan_wait_ticket": (C.c_int, [C.c_void_p, C.c_uint64]),
    "speqg_scan_last_ticket": (C.c_int, [C.c_void_p, u64p]),
    "speqg_scan_flush": (C.c_int, [C.c_void_p]),
    "speqg_scan_fastq_rejected": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, u64p]),
    "speqg_scan_submit_device": (C.c_int, [C.c_void_p, C.POINTER(Batch)]),
    "speqg_scan_sync": (C.c_int, [C.c_void_p]),
    "speqg_scan_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(Counts)]),
    "speqg_scan_reset": (C.c_int, [C.c_void_p]),
    "speqg_scan_stats": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(Counts)]),
    "speqg_scan_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "speqg_scan_fusion_counts": (C.c_int, [C.c_void_p, C.c_void_p]),
    "speqg_scan_fusion_sets": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, u64p, u64p]),
    "speqg_pack_device": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.c_int, C.POINTER(Batch), C.c_void_p]),
    "speqg_probe_random_sectors": (C.c_int, [C.c_int, C.c_uint64, C.POINTER(C.c_double)]),
    "speqg_scan_device_counters": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_void_p),
                                             C.POINTER(C.c_uint64)]),
    "speqg_scan_close": (None, [C.c_void_p]),
    "speqg_batch_to_device": (C.c_int, [C.c_int, C.POINTER(Batch), C.POINTER(Batch)]),
    "speqg_batch_free_device": (None, [C.c_int, C.POINTER(Batch)]),
    "speqg_backcheck": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p]),
    "speqg_refine": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "speqg_scan_multiplicity_count": (C.c_int, [C.c_void_p, u64p]),
    "speqg_refine_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "speqg_unique_to_percent": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
}


class SpeqError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[{code}] {message}")
        self.code = code
        self.message = message


_lib = None


def lib() -> C.CDLL:
    """Load libspeq_cuda.so (built by __graft_entry__.build() / speq_b200/build.py).  Fails loudly when missing."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing: run `python -m speq_b200.build` (nvcc, sm_100a). "
                              "There is no CPU fallback for the scan path.")
        l = C.CDLL(os.fspath(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            f = getattr(l, name)
            f.restype = res
            f.argtypes = args
        _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != OK:
        raise SpeqError(rc, lib().speqg_last_error().decode(errors="replace"))


def ptr(a) -> int | None:
    """Raw address of a numpy array / torch tensor / int (None -> NULL)."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    return a.data_ptr()   # torch tensor
