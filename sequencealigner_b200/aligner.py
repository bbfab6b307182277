"""Host-side mirror of the reference's interface for the all-vs-all scoring path.

`AlignConfig` carries what the reference CLI passes with `-a/-m/-p/-s/-e/-f`
(reference: src/bio/align.c:111-202, src/bio/matrices.c:44-57, src/bio/filter.c:91-117)
and applies the same validation rules; `Aligner` drives the C ABI
(include/seqalign_b200.h) the way `cuda_align` drives the reference backend
(src/interface/seqalign_cuda.c:95-296): upload the sequence store, score every
i<j pair, fill the similarity matrix in the layout `output_fill` produces
(src/io/output.c:68-84).

Everything numeric happens in libseqalign_b200.so on the GPU.  There is no CPU
fallback here: a missing library or device raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import capi
from .seqstore import SeqStore

METHOD_ALIASES = {  # align.c:87-96, matched case-insensitively
    "needleman-wunsch": capi.SAB_NW, "nw": capi.SAB_NW,
    "gotoh": capi.SAB_GA, "ga": capi.SAB_GA,
    "smith-waterman": capi.SAB_SW, "sw": capi.SAB_SW,
}
METHOD_NAMES = {capi.SAB_NW: "nw", capi.SAB_GA: "ga", capi.SAB_SW: "sw"}


def matrix_names() -> list[str]:
    lib = capi.load()
    return [lib.sab_matrix_name(k).decode() for k in range(lib.sab_matrix_count())]


def matrix_lookup(name: str) -> tuple[np.ndarray, np.ndarray]:
    """SEQ_LUT[128], SUB_MAT[24][24] of table `name` (case-insensitive, matrices.c:47)."""
    lib = capi.load()
    lut = np.zeros(128, dtype=np.int32)
    sub = np.zeros(576, dtype=np.int32)
    rc = lib.sab_matrix_lookup(name.encode(), lut.ctypes.data_as(C.POINTER(C.c_int32)),
                               sub.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc != capi.SAB_OK:
        raise ValueError(f"Unknown substitution matrix: {name}")
    return lut, sub.reshape(24, 24)


def filter_resolve(num: int, pairs: np.ndarray) -> np.ndarray:
    """Phase B of the filter (host only): the sequential recurrence of filter.c:35-55 over the
    concatenated match lists of all shards.  Returns the lost[] mask."""
    lib = capi.load()
    pairs = np.ascontiguousarray(pairs, dtype=np.uint64)
    lost = np.zeros(num, dtype=np.uint8)
    rc = lib.sab_filter_resolve(num, pairs.ctypes.data, pairs.size, lost.ctypes.data)
    if rc < 0:
        raise capi.SabError(rc, "sab_filter_resolve")
    return lost.astype(bool)


def plan_strips(store: SeqStore, nshards: int = 1, panel: int = 0) -> list[tuple[int, int, int]]:
    """Host-only row-strip plan: (row0, rows, owner shard) per strip -- which rank of a
    multi-GPU job fills which rows of the matrix (no device needed)."""
    lib = capi.load()
    meta = np.ascontiguousarray(store.meta)
    mp = meta.ctypes.data_as(C.POINTER(capi.SabMeta))
    n = lib.sab_plan_strips(mp, store.num, panel, nshards, 0, None, None, None)
    if n < 0:
        raise capi.SabError(n, "sab_plan_strips")
    a, b, c = (np.zeros(n, np.int32) for _ in range(3))
    p32 = C.POINTER(C.c_int32)
    lib.sab_plan_strips(mp, store.num, panel, nshards, n, a.ctypes.data_as(p32),
                        b.ctypes.data_as(p32), c.ctypes.data_as(p32))
    return list(zip(a.tolist(), b.tolist(), c.tolist()))


@dataclass
class AlignConfig:
    """CLI-level configuration.  Gap values are given as the CLI takes them
    (non-negative) and stored negated, like parse_gap_value (align.c:127-128)."""
    method: str = "nw"
    matrix: str = "blosum62"
    gap_pen: int | None = None   # -p, linear methods only
    gap_open: int | None = None  # -s, affine methods only
    gap_ext: int | None = None   # -e, affine methods only
    filter: float = 0.0          # -f

    def resolve(self) -> dict:
        key = self.method.lower()
        if key not in METHOD_ALIASES:
            raise ValueError(f"Unknown alignment method: {self.method}")
        method = METHOD_ALIASES[key]
        affine = method != capi.SAB_NW
        if affine:
            if self.gap_pen is not None:
                raise ValueError("-p, --gap-penalty is invalid for affine gap methods")
            if self.gap_open is None or self.gap_ext is None:
                raise ValueError("-s, --gap-open and -e, --gap-extend are required")
        else:
            if self.gap_open is not None or self.gap_ext is not None:
                raise ValueError("-s/-e are invalid for linear gap methods")
            if self.gap_pen is None:
                raise ValueError("-p, --gap-penalty is required")
        for v in (self.gap_pen, self.gap_open, self.gap_ext):
            if v is not None and (v < 0 or v > 2**31 - 1):
                raise ValueError("gap values are non-negative integers")
        if not 0.0 <= self.filter <= 1.0:
            raise ValueError("filter threshold must be in [0.0, 1.0]")
        gap_pen = -(self.gap_pen or 0)
        gap_open = -(self.gap_open or 0)
        gap_ext = -(self.gap_ext or 0)
        # validate_ga (ga.c:70-88): equal open/extend degenerates to linear NW
        if method == capi.SAB_GA and gap_open == gap_ext:
            method, gap_pen = capi.SAB_NW, gap_open
        lut, sub = matrix_lookup(self.matrix)
        return dict(method=method, gap_pen=gap_pen, gap_open=gap_open, gap_ext=gap_ext,
                    seq_lut=lut, sub_mat=sub)


class Aligner:
    def __init__(self, config: AlignConfig | dict, device: int = 0, panel: int = 0,
                 flags: int = 0, shards_hint: int = 0):
        self._lib = capi.load()
        r = config.resolve() if isinstance(config, AlignConfig) else dict(config)
        self.resolved = r
        cfg = capi.SabConfig()
        cfg.method = int(r["method"])
        cfg.gap_pen, cfg.gap_open, cfg.gap_ext = int(r["gap_pen"]), int(r["gap_open"]), int(r["gap_ext"])
        cfg.seq_lut[:] = [int(v) for v in np.asarray(r["seq_lut"]).reshape(128)]
        cfg.sub_mat[:] = [int(v) for v in np.asarray(r["sub_mat"]).reshape(576)]
        cfg.device, cfg.panel, cfg.flags, cfg.shards_hint = device, panel, flags, shards_hint
        self._ctx = C.c_void_p()
        rc = self._lib.sab_create(C.byref(self._ctx), C.byref(cfg))
        if rc != capi.SAB_OK:
            msg = self._lib.sab_last_error(self._ctx).decode() if self._ctx else "sab_create failed"
            if self._ctx:
                self._lib.sab_destroy(self._ctx)
                self._ctx = C.c_void_p()
            raise capi.SabError(rc, msg)
        self.num = 0

    def _check(self, rc: int):
        if rc != capi.SAB_OK:
            raise capi.SabError(rc, self._lib.sab_last_error(self._ctx).decode())

    def upload(self, store: SeqStore) -> None:
        seqs = np.ascontiguousarray(store.seqs, dtype=np.uint8)
        meta = np.ascontiguousarray(store.meta)
        self._check(self._lib.sab_upload(self._ctx, seqs.ctypes.data,
                                         meta.ctypes.data_as(C.POINTER(capi.SabMeta)),
                                         store.num, store.max))
        self.num = store.num

    def align(self, triangular: bool = True, out: np.ndarray | None = None, write: bool = True,
              shard: int = 0, nshards: int = 1, progress=None) -> np.ndarray | None:
        """Scores every i<j pair.  Returns the matrix in the reference's layout:
        packed triangular int32[n(n-1)/2] (matrix[j(j-1)/2+i]) or full int32[n,n]
        (both triangles, diagonal 0).  write=False mirrors -W / --no-write."""
        n = self.num
        if write:
            if out is None:
                out = np.zeros(n * (n - 1) // 2 if triangular else (n, n), dtype=np.int32)
            want = n * (n - 1) // 2 if triangular else n * n
            if out.dtype != np.int32 or out.size != want or not out.flags.c_contiguous:
                raise ValueError("out must be a C-contiguous int32 array of the layout's size")
            ptr = out.ctypes.data
        else:
            out, ptr = None, None
        cb = capi.PROGRESS_FN(progress) if progress else None
        self._check(self._lib.sab_align_shard(self._ctx, shard, nshards, ptr, int(triangular),
                                              C.cast(cb, C.c_void_p) if cb else None, None))
        return out

    def strips(self, nshards: int = 1) -> list[tuple[int, int, int]]:
        """(row0, rows, owner shard) of every row strip."""
        res = []
        for k in range(self._lib.sab_strip_count(self._ctx)):
            a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
            self._check(self._lib.sab_strip_info(self._ctx, k, nshards, C.byref(a), C.byref(b),
                                                 C.byref(c)))
            res.append((a.value, b.value, c.value))
        return res

    def stats(self) -> dict:
        st = capi.SabStats()
        self._check(self._lib.sab_get_stats(self._ctx, C.byref(st)))
        return st.as_dict()

    def filter(self, store: SeqStore, threshold: float) -> np.ndarray:
        """filter() decisions (filter.c:14-62, sequential semantics): lost[j] mask."""
        lost = np.zeros(store.num, dtype=np.uint8)
        seqs = np.ascontiguousarray(store.seqs, dtype=np.uint8)
        meta = np.ascontiguousarray(store.meta)
        rc = self._lib.sab_filter(self._ctx, seqs.ctypes.data,
                                  meta.ctypes.data_as(C.POINTER(capi.SabMeta)), store.num,
                                  C.c_float(threshold), lost.ctypes.data)
        if rc < 0:
            self._check(rc)
        return lost.astype(bool)

    def filter_matches(self, store: SeqStore, threshold: float, shard: int = 0, nshards: int = 1,
                       cap: int = 1 << 22) -> np.ndarray:
        """Phase A of the filter for one shard of the j-tiles: uint64 (j << 32 | i) of every pair
        whose prefix identity reaches the threshold.  Raises SAB_ERR_OVERFLOW beyond `cap`."""
        pairs = np.zeros(cap, dtype=np.uint64)
        count = C.c_int64()
        seqs = np.ascontiguousarray(store.seqs, dtype=np.uint8)
        meta = np.ascontiguousarray(store.meta)
        self._check(self._lib.sab_filter_matches(
            self._ctx, seqs.ctypes.data, meta.ctypes.data_as(C.POINTER(capi.SabMeta)), store.num,
            C.c_float(threshold), shard, nshards, pairs.ctypes.data, cap, C.byref(count)))
        return pairs[:count.value].copy()

    def close(self) -> None:
        if self._ctx:
            self._lib.sab_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiAligner:
    """All GPUs of the box from one process (sab_multi_*): what the reference-signature shim
    drives.  ndev = 0 uses every usable device."""

    def __init__(self, config: AlignConfig | dict, ndev: int = 0, panel: int = 0, flags: int = 0):
        self._lib = capi.load()
        r = config.resolve() if isinstance(config, AlignConfig) else dict(config)
        cfg = capi.SabConfig()
        cfg.method = int(r["method"])
        cfg.gap_pen, cfg.gap_open, cfg.gap_ext = int(r["gap_pen"]), int(r["gap_open"]), int(r["gap_ext"])
        cfg.seq_lut[:] = [int(v) for v in np.asarray(r["seq_lut"]).reshape(128)]
        cfg.sub_mat[:] = [int(v) for v in np.asarray(r["sub_mat"]).reshape(576)]
        cfg.panel, cfg.flags = panel, flags
        self._m = C.c_void_p()
        rc = self._lib.sab_multi_create(C.byref(self._m), C.byref(cfg), ndev)
        if rc != capi.SAB_OK:
            msg = self._lib.sab_multi_last_error(self._m).decode() if self._m else "sab_multi_create"
            self.close()
            raise capi.SabError(rc, msg)
        self.num = 0

    @property
    def ndev(self) -> int:
        """devices holding the current set (after upload), else the devices requested/available"""
        return self._lib.sab_multi_device_count(self._m)

    def _check(self, rc: int):
        if rc != capi.SAB_OK:
            raise capi.SabError(rc, self._lib.sab_multi_last_error(self._m).decode())

    def upload(self, store: SeqStore) -> None:
        seqs = np.ascontiguousarray(store.seqs, dtype=np.uint8)
        meta = np.ascontiguousarray(store.meta)
        self._check(self._lib.sab_multi_upload(self._m, seqs.ctypes.data,
                                               meta.ctypes.data_as(C.POINTER(capi.SabMeta)),
                                               store.num, store.max))
        self.num = store.num

    def align(self, triangular: bool = True, out: np.ndarray | None = None, write: bool = True,
              progress=None) -> np.ndarray | None:
        n = self.num
        if write:
            if out is None:
                out = np.zeros(n * (n - 1) // 2 if triangular else (n, n), dtype=np.int32)
            ptr = out.ctypes.data
        else:
            out, ptr = None, None
        cb = capi.PROGRESS_FN(progress) if progress else None
        self._check(self._lib.sab_multi_align(self._m, ptr, int(triangular),
                                              C.cast(cb, C.c_void_p) if cb else None, None))
        return out

    def filter(self, store: SeqStore, threshold: float) -> np.ndarray:
        """filter() decisions with phase A spread over the devices (sab_multi_filter)."""
        lost = np.zeros(store.num, dtype=np.uint8)
        seqs = np.ascontiguousarray(store.seqs, dtype=np.uint8)
        meta = np.ascontiguousarray(store.meta)
        rc = self._lib.sab_multi_filter(self._m, seqs.ctypes.data,
                                        meta.ctypes.data_as(C.POINTER(capi.SabMeta)), store.num,
                                        C.c_float(threshold), lost.ctypes.data)
        if rc < 0:
            self._check(rc)
        return lost.astype(bool)

    def stats(self) -> dict:
        st = capi.SabStats()
        self._check(self._lib.sab_multi_get_stats(self._m, C.byref(st)))
        return st.as_dict()

    def close(self) -> None:
        if self._m:
            self._lib.sab_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def align_all(store: SeqStore, config: AlignConfig, triangular: bool = True, device: int = 0,
              **kw) -> tuple[np.ndarray, SeqStore]:
    """input_load's filter step + cuda_align in one call (main.c:22-32): returns the
    matrix and the (possibly filtered) store it indexes."""
    with Aligner(config, device=device, **kw) as al:
        if config.filter > 0.0:
            lost = al.filter(store, config.filter)
            store = store.subset(~lost)
            if store.num < 2:
                raise ValueError(f"Not enough sequences: {store.num} (min: 2)")
        al.upload(store)
        return al.align(triangular=triangular), store
