"""GPU: sab_sw_pairs_noalloc -- the entry point that allocates nothing (static device scratch; no counterpart
in the reference) -- against the existing entry point
(sab_upload + sab_align, Smith-Waterman) through the C ABI, bit for bit, on two calls in a row.

Collected last on purpose (file name).  The device buffers the entry point
takes are allocated by the TEST (plain cudaMalloc through ctypes on the CUDA runtime the library itself is
linked against): the entry point allocates nothing."""
import ctypes as C

import numpy as np
import pytest

from sequencealigner_b200 import AlignConfig, Aligner, capi, synth

H2D, D2H = 1, 2


def cudart():
    """The libcudart the library is linked against (already mapped once the library is loaded)."""
    capi.load()
    path = next((l.split()[-1] for l in open("/proc/self/maps") if "libcudart.so" in l), "libcudart.so.12")
    rt = C.CDLL(path)
    rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    rt.cudaMemset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
    rt.cudaFree.argtypes = [C.c_void_p]
    rt.cudaGetErrorString.restype = C.c_char_p
    return rt


class DeviceBuffers:
    def __init__(self, rt):
        self.rt, self.ptrs = rt, []

    def put(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.empty(arr.nbytes)
        assert self.rt.cudaMemcpy(p, arr.ctypes.data, arr.nbytes, H2D) == 0
        return p

    def empty(self, nbytes):
        p = C.c_void_p()
        rc = self.rt.cudaMalloc(C.byref(p), max(nbytes, 4))
        assert rc == 0, self.rt.cudaGetErrorString(rc)
        self.ptrs.append(p)
        return p

    def get(self, p, n, dtype):
        out = np.empty(n, dtype=dtype)
        rc = self.rt.cudaMemcpy(out.ctypes.data, p, out.nbytes, D2H)  # blocking: waits for the default stream
        assert rc == 0, self.rt.cudaGetErrorString(rc)
        return out

    def free(self):
        for p in self.ptrs:
            self.rt.cudaFree(p)


def sab_config(cfg: AlignConfig) -> capi.SabConfig:
    r = cfg.resolve()
    c = capi.SabConfig()
    c.method, c.gap_pen, c.gap_open, c.gap_ext = (int(r["method"]), int(r["gap_pen"]), int(r["gap_open"]),
                                                  int(r["gap_ext"]))
    c.seq_lut[:] = [int(v) for v in np.asarray(r["seq_lut"]).reshape(128)]
    c.sub_mat[:] = [int(v) for v in np.asarray(r["sub_mat"]).reshape(576)]
    return c


@pytest.mark.gpu
def test_no_alloc_entry_point_equals_sab_align_on_two_calls_in_a_row():
    lib = capi.load()
    rt = cudart()
    store, _ = synth.make("c5", n=160)  # proteins, len 50..1000: ragged, rows of very different lengths
    cfg = AlignConfig("sw", "blosum62", gap_open=10, gap_ext=1)
    with Aligner(cfg, device=0) as al:  # the existing entry point
        al.upload(store)
        want = al.align(triangular=True)
    n = store.num
    rng = np.random.default_rng(11)
    sc = sab_config(cfg)
    err = C.create_string_buffer(256)
    dev = DeviceBuffers(rt)
    try:
        d_seqs, d_meta = dev.put(store.seqs), dev.put(store.meta)
        # set-up call (count == 0): looks the scratch address up, launches nothing
        assert lib.sab_sw_pairs_noalloc(C.byref(sc), None, None, n, store.max, None, None, 0, None, None,
                                        err, len(err)) == 0, err.value
        first = None
        for call, count in enumerate((3000, 12720)):  # second call: every pair, each in a random order (i, j) / (j, i)
            if call == 0:
                i = rng.integers(0, n, count).astype(np.int32)
                j = ((i + rng.integers(1, n, count)) % n).astype(np.int32)
            else:
                jj, ii = np.tril_indices(n, -1)
                assert ii.size == count
                flip = rng.random(count) < 0.5
                i = np.where(flip, jj, ii).astype(np.int32)
                j = np.where(flip, ii, jj).astype(np.int32)
            d_i, d_j, d_out = dev.put(i), dev.put(j), dev.empty(4 * count)
            assert rt.cudaMemset(d_out, 0x7f, 4 * count) == 0
            rc = lib.sab_sw_pairs_noalloc(C.byref(sc), d_seqs, d_meta, n, store.max, d_i, d_j, count, d_out,
                                          None, err, len(err))
            assert rc == 0, err.value
            got = dev.get(d_out, count, np.int32)
            lo, hi = np.minimum(i, j).astype(np.int64), np.maximum(i, j).astype(np.int64)
            expect = want[hi * (hi - 1) // 2 + lo]
            bad = np.flatnonzero(got != expect)
            assert bad.size == 0, (call, bad[:5], got[bad[:5]], expect[bad[:5]])  # integers: bit-exact
            if call == 0:
                first = (d_i, d_j, got)
        # the first list again, after the scratch was used by a larger call: same answer
        d_out = dev.empty(4 * 3000)
        assert lib.sab_sw_pairs_noalloc(C.byref(sc), d_seqs, d_meta, n, store.max, first[0], first[1], 3000,
                                        d_out, None, err, len(err)) == 0, err.value
        assert np.array_equal(dev.get(d_out, 3000, np.int32), first[2])

        # limits and argument errors: rejected with a message, nothing launched
        rc = lib.sab_sw_pairs_noalloc(C.byref(sc), d_seqs, d_meta, n, capi.SAB_NOALLOC_MAX_LEN + 1, first[0],
                                      first[1], 10, d_out, None, err, len(err))
        assert rc == -1 and b"SAB_NOALLOC_MAX_LEN" in err.value
        nw = sab_config(AlignConfig("nw", "blosum62", gap_pen=4))
        assert lib.sab_sw_pairs_noalloc(C.byref(nw), d_seqs, d_meta, n, store.max, first[0], first[1], 10, d_out,
                                        None, err, len(err)) == -1 and b"Smith-Waterman only" in err.value
        host = np.zeros(10, np.int32)  # a HOST buffer where a device buffer belongs
        assert lib.sab_sw_pairs_noalloc(C.byref(sc), d_seqs, d_meta, n, store.max, first[0], first[1], 10,
                                        host.ctypes.data, None, err, len(err)) == -1 and b"DEVICE" in err.value
        # a bad pair is flagged in its own slot, the others are scored
        i = np.array([0, 5, n, 3], np.int32)
        j = np.array([1, 5, 0, 2], np.int32)
        d_i, d_j = dev.put(i), dev.put(j)
        assert lib.sab_sw_pairs_noalloc(C.byref(sc), d_seqs, d_meta, n, store.max, d_i, d_j, 4, d_out, None,
                                        err, len(err)) == 0, err.value
        got = dev.get(d_out, 4, np.int32)
        assert got.tolist() == [int(want[0]), -2**31, -2**31, int(want[3 * 2 // 2 + 2])]
    finally:
        dev.free()


def test_no_alloc_entry_point_is_declared_and_rejects_bad_arguments_without_a_device():
    """CPU: argument checks come before any CUDA call."""
    lib = capi.load()
    err = C.create_string_buffer(256)
    sc = sab_config(AlignConfig("sw", "blosum62", gap_open=10, gap_ext=1))
    assert lib.sab_sw_pairs_noalloc(None, None, None, 0, 1, None, None, 0, None, None, err, len(err)) == -1
    assert lib.sab_sw_pairs_noalloc(C.byref(sc), None, None, 2, 5000, None, None, 1, None, None, err,
                                    len(err)) == -1 and b"4095" in err.value
    assert capi.SAB_NOALLOC_SCRATCH_BYTES == 16 << 20
    hdr = (capi.LIB.parent.parent / "include" / "seqalign_b200.h").read_text()
    assert "#define SAB_NOALLOC_MAX_LEN 4095" in hdr and "(16u << 20)" in hdr
    src = (capi.LIB.parent / "csrc" / "sab_noalloc.cu").read_text()
    code = "\n".join(l.split("//")[0] for l in src.splitlines())
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMemcpy", "new ", "std::vector", "malloc("):
        assert banned not in code, banned
    assert "static __device__ __align__(256)" in code and "cudaGetSymbolAddress" in code
