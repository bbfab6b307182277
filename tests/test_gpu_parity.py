"""GPU parity tests proper: the CUDA path, called through the C ABI
(include/seqalign_b200.h via ctypes), against
  * the committed golden fixtures (outputs of the unmodified reference CPU build),
  * the CPU oracle on the same seeded inputs,
  * size-independent properties at larger sizes.
Bit-exact: these are int32 scores and filter decisions."""
import numpy as np
import pytest

from conftest import GOLDEN, cli_to_config, golden_cases, load_golden, parse_fasta_bytes
from sequencealigner_b200 import AlignConfig, Aligner, SeqStore, align_all, capi, synth

pytestmark = pytest.mark.gpu

METHODS = [("nw", dict(gap_pen=4)), ("ga", dict(gap_open=10, gap_ext=1)),
           ("sw", dict(gap_open=10, gap_ext=1))]


def gpu_scores(cfg, store, triangular=True, **kw):
    with Aligner(cfg, **kw) as al:
        al.upload(store)
        out = al.align(triangular=triangular)
        return out, al.stats()


@pytest.mark.parametrize("name", golden_cases())
def test_golden(name):
    g = load_golden(name)
    cfg = cli_to_config(g["args"])
    mat, kept = align_all(SeqStore.from_sequences(g["seqs"]), cfg, triangular=g["triangular"])
    assert [kept.sequence(k) for k in range(kept.num)] == g["kept"]
    assert np.array_equal(mat, g["matrix"])


@pytest.mark.parametrize("flags", [0, capi.SAB_FLAG_FORCE_INT32])
def test_golden_all_tables(flags):
    z = np.load(GOLDEN / "all_tables.npz")
    stores = {k: SeqStore.from_sequences(parse_fasta_bytes(z[k])) for k in ("protein", "nucleo")}
    for key in z.files:
        if "/" not in key:
            continue
        name, method = key.split("/")
        cfg = AlignConfig(method=method, matrix=name, gap_pen=3) if method == "nw" else \
            AlignConfig(method=method, matrix=name, gap_open=9, gap_ext=2)
        store = stores["nucleo" if name in ("dnafull", "nuc44") else "protein"]
        mat, _ = gpu_scores(cfg, store, triangular=False, flags=flags)
        assert np.array_equal(mat, z[key]), key


@pytest.mark.parametrize("method,kw", METHODS)
@pytest.mark.parametrize("flags", [0, capi.SAB_FLAG_FORCE_INT32, capi.SAB_FLAG_STRIP16,
                                   capi.SAB_FLAG_STRIP32, capi.SAB_FLAG_SERIAL_WIDTHS,
                                   capi.SAB_FLAG_STRIP16 | capi.SAB_FLAG_FORCE_INT32])
def test_vs_oracle_mixed_lengths(oracle_mod, method, kw, flags):
    """ragged lengths 1..420 across several panels, strips and register windows"""
    rng = np.random.default_rng(5)
    lens = np.concatenate([rng.integers(1, 420, size=560), [1, 1, 2, 31, 32, 33, 127, 128, 129,
                                                              255, 256, 257, 384, 385, 419]])
    rng.shuffle(lens)
    store = SeqStore.from_codes(rng.integers(0, 20, size=int(lens.sum())), lens, synth.PROTEIN20)
    cfg = AlignConfig(method, "blosum62", **kw)
    want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=True)
    got, st = gpu_scores(cfg, store, triangular=True, panel=256, flags=flags)
    assert st["pairs"] == store.pairs() and st["cells"] == store.cells()
    assert (st["tiles_int32"] > 0) == bool(flags & capi.SAB_FLAG_FORCE_INT32)
    bad = np.flatnonzero(got != want)
    assert bad.size == 0, (bad[:10], got[bad[:10]], want[bad[:10]])


@pytest.mark.parametrize("method,kw", METHODS)
@pytest.mark.parametrize("flags", [0, capi.SAB_FLAG_STRIP32, capi.SAB_FLAG_FORCE_INT32])
def test_vs_oracle_short_sequences_large_work_items(oracle_mod, method, kw, flags):
    """lengths 1..40: one or two strips per subject pair, so a work item takes 64..128 subjects
    (adaptive chunk, sab_capi.cu) and many pairs share the SW running-maximum slots"""
    rng = np.random.default_rng(11)
    lens = rng.integers(1, 41, size=1500)
    store = SeqStore.from_codes(rng.integers(0, 20, size=int(lens.sum())), lens, synth.PROTEIN20)
    cfg = AlignConfig(method, "blosum62", **kw)
    want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=True)
    got, st = gpu_scores(cfg, store, triangular=True, panel=512, flags=flags)
    assert st["pairs"] == store.pairs() and st["cells"] == store.cells()
    bad = np.flatnonzero(got != want)
    assert bad.size == 0, (bad[:10], got[bad[:10]], want[bad[:10]])


@pytest.mark.parametrize("config", ["c1", "c2", "c3", "c4"])
def test_vs_oracle_baseline_configs_scaled(oracle_mod, config):
    """BASELINE.json configs at a size the oracle finishes in seconds"""
    n = {"c1": 700, "c2": 300, "c3": 500, "c4": 900}[config]
    store, c = synth.make(config, n=n)
    cfg = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"),
                      gap_open=c.get("gap_open"), gap_ext=c.get("gap_ext"))
    for tri in (True, False):
        want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=tri)
        got, _ = gpu_scores(cfg, store, triangular=tri)
        assert np.array_equal(got, want)


def test_full_layout_matches_triangular_and_diagonal_untouched():
    store, c = synth.make("c1", n=1000)
    cfg = AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1)
    with Aligner(cfg, panel=256) as al:
        al.upload(store)
        tri = al.align(triangular=True)
        full = np.full((store.num, store.num), 7, dtype=np.int32)
        al.align(triangular=False, out=full)
    assert (np.diag(full) == 7).all()  # output_fill never writes the diagonal (output.c:76-83)
    assert (np.tril(full, -1) == np.triu(full, 1).T).all()
    iu = np.tril_indices(store.num, -1)
    assert np.array_equal(full[iu], tri)  # tril order is (j, i<j) row-major == packed layout


def test_no_write_mode_and_stats():
    store, _ = synth.make("c3", n=600)
    cfg = AlignConfig("sw", "nuc44", gap_open=10, gap_ext=1)
    with Aligner(cfg) as al:
        al.upload(store)
        assert al.align(write=False) is None  # -W / --no-write
        st = al.stats()
    assert st["pairs"] == store.pairs() and st["cells"] == store.cells()
    assert st["d2h_bytes"] == 0 and st["kernel_ms"] > 0 and st["launches"] >= 1


@pytest.mark.parametrize("method,kw", METHODS)
def test_reupload_on_one_context_grow_and_shrink(oracle_mod, method, kw):
    """sab_upload keeps its device buffers and re-uses them when they are large enough: results
    of small -> large -> small uploads on ONE context must equal fresh-context results"""
    small, _ = synth.make("c4", n=300)
    large, _ = synth.make("c2", n=260)
    cfg = AlignConfig(method, "blosum62", **kw)
    want_small, _ = oracle_mod.align_all(cfg.resolve(), small, triangular=True)
    want_large, _ = oracle_mod.align_all(cfg.resolve(), large, triangular=True)
    with Aligner(cfg) as al:
        for store, want in ((small, want_small), (large, want_large), (small, want_small),
                            (large, want_large)):
            al.upload(store)
            got = al.align(triangular=True)
            assert np.array_equal(got, want)
            assert al.stats()["pairs"] == store.pairs()


def test_sw_int16_overflow_rerun(oracle_mod):
    """C5-ovf: poly-W under pam500 exceeds int16 -> a-posteriori int32 re-run list"""
    store, c = synth.make_overflow(n=300, planted=6)
    cfg = AlignConfig("sw", "pam500", gap_open=10, gap_ext=1)
    want, _ = oracle_mod.align_all(cfg.resolve(), store)
    assert want.max() == 34000
    got, st = gpu_scores(cfg, store, flags=capi.SAB_FLAG_FORCE_INT16)
    assert st["rerun_pairs"] >= 15
    assert np.array_equal(got, want)
    got, st = gpu_scores(cfg, store)  # default routing
    assert np.array_equal(got, want)


@pytest.mark.parametrize("method,kw", [("nw", dict(gap_pen=40)), ("ga", dict(gap_open=60, gap_ext=25))])
def test_global_methods_route_wide_ranges_to_int32(oracle_mod, method, kw):
    store, _ = synth.make_overflow(n=200, planted=3)
    cfg = AlignConfig(method, "pam500", **kw)
    want, _ = oracle_mod.align_all(cfg.resolve(), store)
    assert want.min() < -32768 or want.max() > 32767
    got, st = gpu_scores(cfg, store)
    assert st["tiles_int32"] == st["tiles"]
    assert np.array_equal(got, want)


def test_two_sequences_and_tiny_sets(oracle_mod):
    for seqs in (["A", "W"], ["ACDEFGHIKLMNPQRSTVWY", "W"], ["M" * 700, "MK" * 300, "K"]):
        store = SeqStore.from_sequences(seqs)
        for method, kw in METHODS:
            cfg = AlignConfig(method, "blosum62", **kw)
            want, _ = oracle_mod.align_all(cfg.resolve(), store)
            got, _ = gpu_scores(cfg, store)
            assert np.array_equal(got, want), (seqs, method)


def test_errors():
    cfg = AlignConfig("nw", "blosum62", gap_pen=4)
    with Aligner(cfg) as al:
        with pytest.raises(capi.SabError):
            al.align()  # before upload
        with pytest.raises(capi.SabError):
            al.upload(SeqStore.from_sequences(["ACD"]))  # SEQ_N_MIN = 2
        with pytest.raises(capi.SabError):
            al.upload(SeqStore.from_sequences(["ACD", "AJC"]))  # J is not in the alphabet
    r = cfg.resolve()
    r["sub_mat"] = r["sub_mat"].copy()
    r["sub_mat"][0, 1] += 1
    with pytest.raises(capi.SabError):
        Aligner(r)  # asymmetric table
    with pytest.raises(capi.SabError):
        Aligner(cfg, device=99)


def test_panel_size_and_sharding_do_not_change_results():
    """idempotence: any panel height / any shard split gives the same matrix"""
    store, c = synth.make("c5", n=1500)
    cfg = AlignConfig("sw", "blosum62", gap_open=10, gap_ext=1)
    base, _ = gpu_scores(cfg, store, panel=128)
    for panel in (256, 512, 0):
        got, _ = gpu_scores(cfg, store, panel=panel)
        assert np.array_equal(got, base)
    merged = np.full_like(base, -1)
    with Aligner(cfg, panel=128, shards_hint=3) as al:
        al.upload(store)
        owners = al.strips(3)
        for shard in range(3):
            part = np.full_like(base, -1)
            al.align(out=part, shard=shard, nshards=3)
            for row0, rows, owner in owners:
                lo, hi = row0 * (row0 - 1) // 2, (row0 + rows) * (row0 + rows - 1) // 2
                if owner == shard:
                    merged[lo:hi] = part[lo:hi]
                else:
                    assert (part[lo:hi] == -1).all()  # other shards' rows untouched
    assert np.array_equal(merged, base)
    # auto panels = equal-cost strips + a geometric tail per shard (panel_bounds, sab_capi.cu)
    merged = np.full_like(base, -1)
    with Aligner(cfg, shards_hint=4) as al:
        al.upload(store)
        owners = al.strips(4)
        heights = [rows for _, rows, _ in owners]
        assert len(owners) >= 8 and len(set(heights)) > 1 and all(h % 32 == 0 for h in heights[:-1])
        for shard in range(4):
            al.align(out=merged, shard=shard, nshards=4)  # the shards fill disjoint rows of ONE matrix
    assert np.array_equal(merged, base)


@pytest.mark.parametrize("thr", [0.9, 0.5, 0.05, 1.0])
def test_filter_decisions(oracle_mod, thr):
    store, c = synth.make("c4", n=5000)
    want = oracle_mod.filter_lost(store, thr)
    with Aligner(AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1)) as al:
        got = al.filter(store, thr)
    assert want.sum() > 0 and np.array_equal(got, want)


def test_filter_ragged_lengths(oracle_mod):
    rng = np.random.default_rng(3)
    abc = np.frombuffer(b"ACDEFGHIK", dtype=np.uint8)
    base = [abc[rng.integers(0, 9, size=int(rng.integers(1, 90)))].tobytes() for _ in range(40)]
    seqs = []
    for k in range(3000):
        s = base[int(rng.integers(0, 40))]
        cut = int(rng.integers(1, len(s) + 1))
        seqs.append(s[:cut] + abc[rng.integers(0, 9, size=int(rng.integers(0, 20)))].tobytes())
    store = SeqStore.from_sequences(seqs)
    with Aligner(AlignConfig("nw", "blosum62", gap_pen=4)) as al:
        for thr in (0.8, 0.3):
            assert np.array_equal(al.filter(store, thr), oracle_mod.filter_lost(store, thr))


def test_full_size_properties_c2_slice():
    """size-independent checks at a size the oracle cannot do in seconds: NW self-consistency
    (score(i,j) <= min over diagonal sums), SW >= 0 and SW >= NW-free bound, checksums stable"""
    store, c = synth.make("c2", n=3000)
    sw, _ = gpu_scores(AlignConfig("sw", "blosum62", gap_open=10, gap_ext=1), store)
    ga, _ = gpu_scores(AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1), store)
    nw, _ = gpu_scores(AlignConfig("nw", "blosum62", gap_pen=10), store)
    assert (sw >= 0).all() and (sw >= ga).all()  # local >= global with the same gaps
    assert (ga >= nw).all()  # affine(open 10, ext 1) never costs more than linear 10
    again, _ = gpu_scores(AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1), store, panel=384)
    assert int(again.astype(np.int64).sum()) == int(ga.astype(np.int64).sum())
    assert np.array_equal(again, ga)


@pytest.mark.parametrize("config,n", [("c2", 10000), ("c3", 20000), ("c4", 30000), ("c5", 3000)])
def test_baseline_sizes_sampled_pairs_and_checksums(oracle_mod, config, n):
    """BASELINE.json workloads at (or near) full size: the whole matrix is computed on the
    GPU; 3000 uniformly drawn pairs plus every pair of 24 whole rows are re-scored by the
    CPU oracle, and the packed-triangular result must equal the full-layout result."""
    store, c = synth.make(config, n=n)
    cfg = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"),
                      gap_open=c.get("gap_open"), gap_ext=c.get("gap_ext"))
    r = cfg.resolve()
    with Aligner(cfg) as al:
        al.upload(store)
        tri = al.align(triangular=True)
        st = al.stats()
    assert st["pairs"] == store.pairs() and st["cells"] == store.cells()
    rng = np.random.default_rng(11)
    js = rng.integers(1, n, size=3000)
    is_ = (rng.random(3000) * js).astype(np.int64)
    for i, j in zip(is_.tolist(), js.tolist()):
        want = oracle_mod.pair(r, store.sequence(j), store.sequence(i))
        assert tri[j * (j - 1) // 2 + i] == want, (i, j)
    # whole rows through the oracle's own all-vs-all driver (columns j of align(), align.c:21-72)
    rows = sorted(set(rng.integers(1, min(n, 2500), size=24).tolist()) | {1})
    for j in rows:
        want, _ = oracle_mod.align_all(r, store.head(j + 1), first_j=j, last_j=j + 1)
        lo = j * (j - 1) // 2
        assert np.array_equal(tri[lo:lo + j], want[lo:lo + j]), j
    if n <= 10000:
        with Aligner(cfg) as al:
            al.upload(store)
            full = al.align(triangular=False)
        iu = np.tril_indices(n, -1)
        assert np.array_equal(full[iu], tri) and (np.diag(full) == 0).all()
        assert np.array_equal(full, full.T)


def test_multi_device_single_process_equals_single_device(oracle_mod):
    """sab_multi_*: all devices of the box from one process (1 device on the test box is fine:
    the same code path, one shard)."""
    from sequencealigner_b200 import MultiAligner
    store, c = synth.make("c5", n=1200)
    cfg = AlignConfig("sw", "blosum62", gap_open=10, gap_ext=1)
    want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=True)
    seen = []
    with MultiAligner(cfg, panel=128) as ma:
        assert ma.ndev >= 1
        ma.upload(store)
        got = ma.align(triangular=True, progress=lambda d, t, u: seen.append((d, t)))
        st = ma.stats()
        full = ma.align(triangular=False)
    assert np.array_equal(got, want)
    assert st["pairs"] == store.pairs() and st["cells"] == store.cells()
    assert seen and seen[-1] == (store.pairs(), store.pairs())
    assert all(a[0] <= b[0] for a, b in zip(seen, seen[1:]))
    assert np.array_equal(full[np.tril_indices(store.num, -1)], want)


@pytest.mark.parametrize("method,kw", METHODS)
def test_long_sequences_beyond_the_reference_cuda_cap(oracle_mod, method, kw):
    """The reference's CUDA backend refuses lengths > 1023 (seqalign_cuda.c:102-105); its CPU
    path does not.  Lengths up to 6000: many wrap-arounds of the strip pipeline, int32 routing
    for NW (the all-gap border leaves int16), int16 for SW."""
    rng = np.random.default_rng(17)
    lens = np.concatenate([rng.integers(1500, 6001, size=20), rng.integers(1, 200, size=12)])
    rng.shuffle(lens)
    store = SeqStore.from_codes(rng.integers(0, 20, size=int(lens.sum())), lens, synth.PROTEIN20)
    cfg = AlignConfig(method, "blosum62", **kw)
    want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=True)
    got, st = gpu_scores(cfg, store, triangular=True)
    assert np.array_equal(got, want)
    if method == "nw":
        assert st["tiles_int32"] > 0


# ---- round 2 ---------------------------------------------------------------------------

@pytest.mark.parametrize("method", ["ga", "sw"])
@pytest.mark.parametrize("gaps", [(2, 5), (0, 1), (3, 3), (0, 0), (1, 30)])
@pytest.mark.parametrize("flags", [0, capi.SAB_FLAG_FORCE_INT32])
def test_gap_extend_dearer_than_open(oracle_mod, method, gaps, flags):
    """-s below -e is legal on the CLI (align.c puts no constraint on them).  Gotoh then
    degenerates to NW with g = open (the border recurrence ga.c:26-37 gives c*o, not o+(c-1)e):
    the library routes it; SW keeps the general cell.  Oracle = the GA/SW recurrences as written."""
    rng = np.random.default_rng(23)
    lens = np.concatenate([rng.integers(1, 200, size=300), [1, 2, 31, 32, 33, 64, 65]])
    store = SeqStore.from_codes(rng.integers(0, 20, size=int(lens.sum())), lens, synth.PROTEIN20)
    cfg = AlignConfig(method, "blosum62", gap_open=gaps[0], gap_ext=gaps[1])
    r = cfg.resolve()
    if gaps[0] != gaps[1]:
        assert r["method"] == (capi.SAB_GA if method == "ga" else capi.SAB_SW)  # oracle sees GA/SW
    want, _ = oracle_mod.align_all(r, store, triangular=True)
    got, _ = gpu_scores(cfg, store, triangular=True, panel=128, flags=flags)
    bad = np.flatnonzero(got != want)
    assert bad.size == 0, (bad[:10], got[bad[:10]], want[bad[:10]])


def test_sw_rerun_list_overflow_resweeps_the_strip_in_int32(oracle_mod):
    """more overflowing pairs than the int16 re-run list holds (cap lowered to 4 by the test
    flag): the strip is swept again by the int32 instantiation instead of failing"""
    store, c = synth.make_overflow(n=300, planted=8)
    cfg = AlignConfig("sw", "pam500", gap_open=10, gap_ext=1)
    want, _ = oracle_mod.align_all(cfg.resolve(), store)
    for tri in (True, False):
        got, st = gpu_scores(cfg, store, triangular=tri, panel=128,
                             flags=capi.SAB_FLAG_TEST_RERUN_CAP | capi.SAB_FLAG_FORCE_INT16)
        assert st["rerun_strips"] >= 1 and st["tiles_int32"] >= 1
        if tri:
            assert np.array_equal(got, want)
        else:
            assert np.array_equal(got[np.tril_indices(store.num, -1)], want)
            assert np.array_equal(got, got.T)


def test_sw_many_identical_high_scoring_sequences(oracle_mod):
    """~1500 identical high-scoring sequences in one panel used to end in
    SAB_ERR_OVERFLOW.  1500 x poly-W(1000) under pam500: every pair scores 34000 > int16, 1.1e6
    of them -> more than 2 % of a strip, so strips are swept again in int32, and once that has
    happened the remaining strips start in int32."""
    seqs = [b"W" * 1000] * 1500
    store = SeqStore.from_sequences(seqs)
    cfg = AlignConfig("sw", "pam500", gap_open=10, gap_ext=1)
    got, st = gpu_scores(cfg, store, triangular=True)
    assert 1 <= st["rerun_strips"] < st["tiles"] and st["rerun_pairs"] == 0
    # every strip ended in int32: swept again after a packed attempt, or started there
    assert st["tiles_int32"] == st["tiles"] - st["rerun_strips"]
    assert (got == 34000).all()
    assert oracle_mod.pair(cfg.resolve(), seqs[5], seqs[0]) == 34000


@pytest.mark.parametrize("thr", [0.9, 0.3])
def test_filter_paths_agree(oracle_mod, thr):
    """match list (default), panel-sequential (flag), list overflow -> fallback (test cap),
    sharded phase A + host resolve, and the multi-device entry point: one answer"""
    from sequencealigner_b200 import MultiAligner
    from sequencealigner_b200.aligner import filter_resolve
    store, c = synth.make("c4", n=6000)
    want = oracle_mod.filter_lost(store, thr)
    cfg = AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1)
    for flags in (0, capi.SAB_FLAG_FILTER_SEQUENTIAL, capi.SAB_FLAG_TEST_FILTER_CAP):
        with Aligner(cfg, flags=flags) as al:
            got = al.filter(store, thr)
            st = al.stats()
        assert np.array_equal(got, want), flags
        assert st["filter_ms"] > 0
        assert (st["filter_matches"] >= 0) == (flags == 0)
    with Aligner(cfg) as al:
        parts = [al.filter_matches(store, thr, shard=k, nshards=3) for k in range(3)]
        whole = al.filter_matches(store, thr)
    assert sorted(np.concatenate(parts).tolist()) == sorted(whole.tolist())
    assert np.array_equal(filter_resolve(store.num, np.concatenate(parts)), want)
    with Aligner(cfg) as al, pytest.raises(capi.SabError):
        al.filter_matches(store, thr, cap=3)
    with MultiAligner(cfg) as ma:
        assert np.array_equal(ma.filter(store, thr), want)
        kept = store.subset(~want)
        ma.upload(kept)  # the same contexts then run the alignment
        got = ma.align()
    ref, _ = oracle_mod.align_all(cfg.resolve(), kept.head(200))
    assert np.array_equal(got[:ref.size], ref)


@pytest.mark.parametrize("flags", [0, capi.SAB_FLAG_ORDERED])
@pytest.mark.parametrize("tri", [True, False])
def test_file_backed_map_shared_sink(tmp_path, flags, tri):
    """f2: the matrix may be a file-backed MAP_SHARED mapping (alloc_mmap, src/system/os.c:103-140):
    plain stores from the library's threads, bit-equal to the in-RAM result, in both layouts, in
    cost order and in ascending-row order"""
    import mmap
    store, c = synth.make("c1", n=1500)
    cfg = AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1)
    n = store.num
    count = n * (n - 1) // 2 if tri else n * n
    path = tmp_path / "sink.bin"
    with open(path, "wb") as f:
        f.truncate(count * 4)
    with Aligner(cfg, panel=256, flags=flags) as al:
        al.upload(store)
        ram = al.align(triangular=tri)
        with open(path, "r+b") as f:
            mm = mmap.mmap(f.fileno(), count * 4, flags=mmap.MAP_SHARED)
            view = np.frombuffer(mm, dtype=np.int32)
            al.align(triangular=tri, out=view if tri else view.reshape(n, n))
            del view
            mm.flush()
            mm.close()
    back = np.fromfile(path, dtype=np.int32)
    assert np.array_equal(back, ram.reshape(-1))


def test_ordered_flag_issues_strips_in_row_order():
    store, c = synth.make("c1", n=1200)
    cfg = AlignConfig("nw", "blosum62", gap_pen=4)
    for flags, check in ((capi.SAB_FLAG_ORDERED, lambda d: d == sorted(d)),
                         (0, lambda d: d != sorted(d))):
        done = []
        with Aligner(cfg, panel=128, flags=flags) as al:
            al.upload(store)
            rows = {0: 0}
            acc = 0
            for row0, nrows, _ in al.strips():
                acc += (row0 + nrows) * (row0 + nrows - 1) // 2 - row0 * (row0 - 1) // 2
            al.align(progress=lambda d, t, u: done.append(d))
        # progress counts pairs: in row order the increments grow (strip k holds more pairs than k-1)
        inc = [b - a for a, b in zip([0] + done, done)]
        assert done[-1] == acc and check(inc[:-1]), (flags, inc)


def test_multi_device_dynamic_cursor_every_strip_exactly_once(oracle_mod):
    """sab_multi_align: the devices pull strips from one shared cursor; whatever the device
    count, every row strip is computed exactly once (checked against the oracle)"""
    from sequencealigner_b200 import MultiAligner
    store, c = synth.make("c1", n=900)
    cfg = AlignConfig("ga", "blosum62", gap_open=10, gap_ext=1)
    want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=True)
    ndev_all = capi.load().sab_device_count()
    for ndev in sorted({1, ndev_all}):
        with MultiAligner(cfg, ndev=ndev, panel=128) as ma:
            ma.upload(store)
            assert ma.ndev == ndev
            out = np.full(want.size, -77, dtype=np.int32)
            ma.align(out=out)
            st = ma.stats()
        assert np.array_equal(out, want)
        assert st["pairs"] == store.pairs() and st["cells"] == store.cells()
        assert st["tiles"] == (store.num + 127) // 128
