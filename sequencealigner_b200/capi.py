"""ctypes binding of the C ABI in include/seqalign_b200.h (no torch types cross it).

The library is the product: if it is missing or cannot be loaded this module
raises -- there is no Python / CPU fallback for the alignment path.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from .build import LIB

SAB_OK = 0
SAB_ERR_PARSE = -7
SAB_NOALLOC_SCRATCH_BYTES = 16 << 20
SAB_NOALLOC_MAX_LEN = 4095
SAB_SRC_NONE, SAB_SRC_FASTA, SAB_SRC_DSV = 0, 1, 2
SAB_NW, SAB_GA, SAB_SW = 0, 1, 2
SAB_FLAG_FORCE_INT32 = 1
SAB_FLAG_FORCE_INT16 = 2
SAB_FLAG_STRIP32 = 8
SAB_FLAG_STRIP16 = 16
SAB_FLAG_ORDERED = 32
SAB_FLAG_TEST_RERUN_CAP = 64
SAB_FLAG_FILTER_SEQUENTIAL = 128
SAB_FLAG_TEST_FILTER_CAP = 256
SAB_FLAG_SERIAL_WIDTHS = 512
SAB_FLAG_PER_CHUNK_WIDTHS = 1024
STATUS = {0: "SAB_OK", -1: "SAB_ERR_ARG", -2: "SAB_ERR_NODEVICE", -3: "SAB_ERR_CUDA",
          -4: "SAB_ERR_NOMEM", -5: "SAB_ERR_STATE", -6: "SAB_ERR_OVERFLOW", -7: "SAB_ERR_PARSE"}


class SabMeta(C.Structure):
    _fields_ = [("off", C.c_int32), ("len", C.c_int32)]


class SabConfig(C.Structure):
    _fields_ = [
        ("method", C.c_int32), ("gap_pen", C.c_int32), ("gap_open", C.c_int32),
        ("gap_ext", C.c_int32), ("seq_lut", C.c_int32 * 128), ("sub_mat", C.c_int32 * 576),
        ("device", C.c_int32), ("panel", C.c_int32), ("flags", C.c_int32),
        ("shards_hint", C.c_int32), ("host_threads", C.c_int32), ("reserved", C.c_int32 * 3),
    ]


class SabStats(C.Structure):
    _fields_ = [
        ("kernel_ms", C.c_double), ("busy_ms", C.c_double), ("device_ms", C.c_double),
        ("total_ms", C.c_double),
        ("upload_ms", C.c_double),
        ("pairs", C.c_int64), ("cells", C.c_int64), ("rerun_pairs", C.c_int64),
        ("launches", C.c_int64), ("tiles", C.c_int64), ("tiles_int32", C.c_int64),
        ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64), ("rerun_strips", C.c_int64),
        ("filter_matches", C.c_int64), ("filter_ms", C.c_double),
        ("width_launches", C.c_int32 * 16),
    ]

    def as_dict(self):
        d = {n: getattr(self, n) for n, _ in self._fields_ if n != "width_launches"}
        d["width_launches"] = {32 - 2 * w: int(v) for w, v in enumerate(self.width_launches) if v}
        return d


PROGRESS_FN = C.CFUNCTYPE(None, C.c_int64, C.c_int64, C.c_void_p)

# every symbol include/seqalign_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "sab_abi_version": (C.c_int, []),
    "sab_device_count": (C.c_int, []),
    "sab_device_ordinal": (C.c_int, [C.c_int]),
    "sab_device_memory": (C.c_int, [C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "sab_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(SabConfig)]),
    "sab_destroy": (None, [C.c_void_p]),
    "sab_last_error": (C.c_char_p, [C.c_void_p]),
    "sab_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(SabMeta), C.c_int32, C.c_int32]),
    "sab_align": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "sab_align_shard": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int,
                                  C.c_void_p, C.c_void_p]),
    "sab_strip_count": (C.c_int, [C.c_void_p]),
    "sab_strip_info": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32),
                                 C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "sab_plan_strips": (C.c_int, [C.POINTER(SabMeta), C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                  C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                  C.POINTER(C.c_int32)]),
    "sab_get_stats": (C.c_int, [C.c_void_p, C.POINTER(SabStats)]),
    "sab_selftest_packing": (C.c_int, [C.POINTER(SabConfig), C.c_void_p, C.POINTER(SabMeta), C.c_int32,
                                       C.POINTER(C.c_int32)]),
    "sab_kernel_image": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t,
                                   C.POINTER(C.c_size_t), C.c_char_p, C.c_size_t]),
    "sab_multi_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(SabConfig), C.c_int32]),
    "sab_multi_destroy": (None, [C.c_void_p]),
    "sab_multi_last_error": (C.c_char_p, [C.c_void_p]),
    "sab_multi_device_count": (C.c_int, [C.c_void_p]),
    "sab_multi_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(SabMeta), C.c_int32, C.c_int32]),
    "sab_multi_align": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "sab_multi_get_stats": (C.c_int, [C.c_void_p, C.POINTER(SabStats)]),
    "sab_multi_filter": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(SabMeta), C.c_int32, C.c_float,
                                   C.c_void_p]),
    "sab_filter_matches": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(SabMeta), C.c_int32, C.c_float,
                                     C.c_int32, C.c_int32, C.c_void_p, C.c_int64,
                                     C.POINTER(C.c_int64)]),
    "sab_filter_resolve": (C.c_int, [C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "sab_filter": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(SabMeta), C.c_int32, C.c_float,
                             C.c_void_p]),
    "sab_matrix_count": (C.c_int, []),
    "sab_matrix_name": (C.c_char_p, [C.c_int]),
    "sab_matrix_lookup": (C.c_int, [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "sab_h5_write": (C.c_int, [C.c_char_p, C.POINTER(C.c_char_p), C.c_size_t, C.c_void_p, C.c_int,
                               C.c_uint, C.c_void_p, C.c_void_p]),
    "sab_h5_write_chunked": (C.c_int, [C.c_char_p, C.POINTER(C.c_char_p), C.c_size_t, C.c_void_p,
                                       C.c_int, C.c_uint, C.c_int, C.c_void_p, C.c_void_p]),
    "sab_h5_chunk_dim": (C.c_int, [C.c_size_t]),
    "sab_source_format": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint8)]),
    "sab_parse_fasta": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_int32), C.c_int32,
                                  C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_char_p, C.c_size_t]),
    "sab_dsv_header": (C.c_int, [C.c_void_p, C.c_size_t, C.c_uint8, C.POINTER(C.c_int32),
                                 C.POINTER(C.c_int32), C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]),
    "sab_parse_dsv": (C.c_int, [C.c_void_p, C.c_size_t, C.c_uint8, C.c_int32, C.c_int,
                                C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_int32),
                                C.POINTER(C.c_int32), C.c_char_p, C.c_size_t]),
    "sab_build_meta": (C.c_int64, [C.c_void_p, C.c_int32, C.POINTER(SabMeta)]),
    "sab_sw_pairs_noalloc": (C.c_int, [C.POINTER(SabConfig), C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                       C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_char_p,
                                       C.c_size_t]),
}

_lib = None


def load(path: Path | None = None) -> C.CDLL:
    """Loads libseqalign_b200.so; raises if it has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path or os.environ.get("SAB_LIB") or LIB)  # SAB_LIB: tuning builds (tools/variants.sh)
    if not p.exists():
        raise RuntimeError(
            f"{p} is missing: build it with `python -m sequencealigner_b200.build` "
            "(there is no CPU fallback for the alignment path)")
    lib = C.CDLL(str(p))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    if lib.sab_abi_version() != 3:
        raise RuntimeError("libseqalign_b200.so ABI version mismatch")
    if path is None:
        _lib = lib
    return lib


class SabError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{STATUS.get(code, code)}: {msg}")
        self.code = code
