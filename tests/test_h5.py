"""sab_h5_write: the HDF5 file of flush_hdf5 (reference: src/io/format/hdf5.c:14-202) written
without libhdf5, checked with the independent structural reader tests/h5min.py.  CPU only."""
import ctypes as C

import numpy as np
import pytest

import h5min
from sequencealigner_b200 import capi


def write(path, seqs, matrix, triangular, compression=0, chunk_dim=-1):
    lib = capi.load()
    arr = (C.c_char_p * len(seqs))(*seqs)
    m = np.ascontiguousarray(matrix, dtype=np.int32)
    rc = lib.sab_h5_write_chunked(str(path).encode(), arr, len(seqs), m.ctypes.data, int(triangular),
                                  compression, chunk_dim, None, None)
    assert rc == capi.SAB_OK
    return rc


def make(n, seed):
    rng = np.random.default_rng(seed)
    abc = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", dtype=np.uint8)
    seqs = [abc[rng.integers(0, 20, size=int(rng.integers(1, 90)))].tobytes() for _ in range(n)]
    tri = rng.integers(-4000, 4000, size=n * (n - 1) // 2).astype(np.int32)
    full = np.zeros((n, n), dtype=np.int32)
    iu = np.tril_indices(n, -1)
    full[iu] = tri  # tril order (j, i<j) row-major == packed layout j(j-1)/2+i (output.c:83)
    full = full + full.T
    return seqs, tri, full


def test_chunk_rule_matches_the_reference_expression():
    lib = capi.load()
    # hdf5.c:68-84: contiguous up to 256, else the largest power of two <= dim in [256, 4096]
    for dim, want in [(2, 0), (256, 0), (257, 256), (511, 256), (512, 512), (1000, 512),
                      (4096, 4096), (5000, 4096), (50000, 4096), (200000, 4096)]:
        assert lib.sab_h5_chunk_dim(dim) == want, dim


@pytest.mark.parametrize("n,compression,triangular", [(2, 0, False), (40, 0, True), (256, 5, True),
                                                      (300, 0, False), (300, 0, True),
                                                      (700, 1, True), (1100, 6, False)])
def test_roundtrip_reference_rule(tmp_path, n, compression, triangular):
    seqs, tri, full = make(n, n)
    p = tmp_path / "out.h5"
    write(p, seqs, tri if triangular else full, triangular, compression)
    got_seqs, got, info = h5min.read_similarity_file(p)
    assert got_seqs == seqs
    assert np.array_equal(got, full) and (np.diag(got) == 0).all()
    cd = capi.load().sab_h5_chunk_dim(n)
    if cd == 0:
        assert info["layout"][0] == "contiguous" and "deflate" not in info  # hdf5.c:69
    else:
        assert info["layout"][0] == "chunked" and info["layout"][2] == (cd, cd, 4)
        assert ("deflate" in info) == (compression > 0)
        if compression:
            assert info["deflate"] == compression


def test_multi_level_chunk_index(tmp_path):
    """16x16 chunks -> 81^2 = 6561 chunks -> 103 leaf nodes, 2 internal levels"""
    n = 1290
    seqs, tri, full = make(n, 5)
    p = tmp_path / "deep.h5"
    write(p, seqs, tri, True, compression=1, chunk_dim=16)
    got_seqs, got, info = h5min.read_similarity_file(p)
    assert got_seqs == seqs and np.array_equal(got, full)
    assert info["nodes"] == 103 + 2 + 1


def test_many_long_strings_span_several_heap_collections(tmp_path):
    rng = np.random.default_rng(9)
    seqs = [bytes(rng.integers(65, 90, size=int(rng.integers(1, 30000))).astype(np.uint8))
            for _ in range(150)]
    full = np.arange(150 * 150, dtype=np.int32).reshape(150, 150)
    p = tmp_path / "s.h5"
    write(p, seqs, full, False)
    got_seqs, got, _ = h5min.read_similarity_file(p)
    assert got_seqs == seqs and np.array_equal(got, full)


def test_bad_arguments():
    lib = capi.load()
    assert lib.sab_h5_write(b"/nonexistent-dir/x.h5", (C.c_char_p * 1)(b"A"), 1,
                            np.zeros(1, np.int32).ctypes.data, 0, 0, None, None) != capi.SAB_OK
    assert lib.sab_h5_write(b"/tmp/x.h5", (C.c_char_p * 1)(b"A"), 1,
                            np.zeros(1, np.int32).ctypes.data, 0, 10, None, None) == -1


def test_reader_walks_a_file_written_by_libhdf5():
    """Pins tests/h5min.py itself: a MATLAB v7.3 .mat file IS an HDF5 file written by libhdf5
    (behind a 512-byte user block).  This one ships with SciPy's test data
    (scipy/io/matlab/tests/data/testhdf5_7.4_GLNX86.mat, BSD licence; 4 KB, committed under
    tests/golden/) and holds one variable: testdouble = 0, pi/4, ..., 2 pi.  The same reader
    code that checks sab_h5_write's files must walk it: superblock v0 + base address, root
    symbol-table entry, group B-tree, symbol-table node, local heap, v1 object header,
    dataspace, datatype, contiguous layout, raw data."""
    import numpy as np
    from conftest import GOLDEN
    import h5min
    f = h5min.File(GOLDEN / "ext_libhdf5_matlab73_testdouble.mat", strict=False)
    links = f.links()
    assert list(links) == ["testdouble"]
    data, info = f.read(links["testdouble"])
    assert info["shape"] == (9, 1) and info["dtype"] == "<f8"
    assert np.allclose(data.ravel(), np.arange(9) * np.pi / 4, rtol=0, atol=1e-15)
    # and the strict mode used for our own files must refuse it (user block, foreign messages)
    import pytest
    with pytest.raises(AssertionError):
        h5min.File(GOLDEN / "ext_libhdf5_matlab73_testdouble.mat")
