"""CPU-only: host logic and the C-ABI surface (no compute calls without a GPU)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from sequencealigner_b200 import AlignConfig, SeqStore, capi, matrix_lookup, matrix_names, synth
from sequencealigner_b200.aligner import plan_strips

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol():
    hdr = (ROOT / "include" / "seqalign_b200.h").read_text()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(sab_[a-z0-9_]+)\s*\(", hdr)) - {"sab_progress_fn"}
    lib = C.CDLL(str(capi.LIB))
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    assert capi.load().sab_abi_version() == 3


def test_struct_sizes_match_header():
    assert C.sizeof(capi.SabMeta) == 8
    assert C.sizeof(capi.SabConfig) == 4 * (4 + 128 + 576 + 4 + 4)
    assert C.sizeof(capi.SabStats) == 8 * 16 + 4 * 16


def test_no_cpu_fallback_without_device():
    lib = capi.load()
    if lib.sab_device_count() > 0:
        pytest.skip("a device is present")
    from sequencealigner_b200 import Aligner
    with pytest.raises(capi.SabError) as e:
        Aligner(AlignConfig("nw", "blosum62", gap_pen=4))
    assert e.value.code == -2  # SAB_ERR_NODEVICE


def test_matrix_tables():
    names = matrix_names()
    assert len(names) == 67 and "blosum62" in names and "nuc44" in names
    lut, sub = matrix_lookup("BLOSUM62")  # case-insensitive, matrices.c:47
    order = "ARNDCQEGHILKMFPSTWYVBZX*"
    assert [lut[ord(c)] for c in order] == list(range(24))
    assert lut[ord("a")] == -1 and lut[ord("J")] == -1
    assert sub[lut[ord("W")], lut[ord("W")]] == 11 and sub[lut[ord("A")], lut[ord("R")]] == -1
    assert (sub == sub.T).all()
    lut, sub = matrix_lookup("nuc44")
    assert [lut[ord(c)] for c in "ATGCSWRYKMBVHDN*"] == list(range(16))
    assert (sub[16:, :] == 0).all()
    with pytest.raises(ValueError):
        matrix_lookup("blosum63")
    for n in names:
        _, s = matrix_lookup(n)
        assert (s == s.T).all() and -23 <= s.min() and s.max() <= 35


def test_align_config_rules():
    r = AlignConfig("Gotoh", "blosum62", gap_open=10, gap_ext=1).resolve()
    assert r["method"] == capi.SAB_GA and r["gap_open"] == -10 and r["gap_ext"] == -1
    r = AlignConfig("ga", "blosum62", gap_open=4, gap_ext=4).resolve()  # validate_ga, ga.c:70-88
    assert r["method"] == capi.SAB_NW and r["gap_pen"] == -4
    with pytest.raises(ValueError):
        AlignConfig("sw", "blosum62", gap_pen=4).resolve()
    with pytest.raises(ValueError):
        AlignConfig("nw", "blosum62", gap_open=4, gap_ext=1).resolve()
    with pytest.raises(ValueError):
        AlignConfig("nw", "blosum62").resolve()
    with pytest.raises(ValueError):
        AlignConfig("xx", "blosum62", gap_pen=1).resolve()
    with pytest.raises(ValueError):
        AlignConfig("nw", "blosum62", gap_pen=1, filter=1.5).resolve()


def test_seqstore_layout_and_counts(tmp_path):
    s = SeqStore.from_sequences(["ACD", "W", "KLMNP"])
    assert s.seqs.tobytes() == b"ACD\0W\0KLMNP\0"
    assert s.meta["off"].tolist() == [0, 4, 6] and s.meta["len"].tolist() == [3, 1, 5]
    assert s.max == 5 and s.pairs() == 3 and s.cells() == 3 * 1 + 3 * 5 + 1 * 5
    with pytest.raises(ValueError):
        SeqStore.from_sequences(["AC", ""])
    p = tmp_path / "a.fasta"
    p.write_bytes(b">x\nac d\nEF\r\n>y\nWW\n")
    from sequencealigner_b200 import read_fasta
    lut, _ = matrix_lookup("blosum62")
    r = read_fasta(p, lut)
    assert [r.sequence(k) for k in range(2)] == [b"ACDEF", b"WW"]
    p.write_bytes(b">x\nAJ\n")
    with pytest.raises(ValueError):
        read_fasta(p, lut)
    store, _ = synth.make("c2", n=50)
    again = SeqStore.from_sequences([store.sequence(k) for k in range(50)])
    assert np.array_equal(store.seqs, again.seqs) and np.array_equal(store.meta, again.meta)
    assert store.head(10).num == 10 and store.head(10).sequence(9) == store.sequence(9)


def test_synth_configs_are_deterministic():
    a, cfg = synth.make("c3", n=64)
    b, _ = synth.make("c3", n=64)
    assert np.array_equal(a.seqs, b.seqs) and set(a.seqs.tobytes()) <= set(b"ACGT\0")
    assert (a.lengths == 150).all() and cfg["method"] == "sw"
    d, cfg = synth.make("c4", n=400)
    assert 30 <= d.lengths.min() and d.lengths.max() <= 60 and cfg["filter"] == 0.9
    o, _ = synth.make_overflow(n=100, planted=4)
    assert sum(o.sequence(k) == b"W" * 1000 for k in range(100)) == 4


def test_strip_plan_covers_rows_once_and_balances():
    """auto panels (sab_capi.cu: panel_bounds): equal-cost row strips, balanced over the shards by
    construction; short sequences (C4) additionally end every shard on a geometric tail of small
    strips, because the write-back of a rank's last strip overlaps with nothing"""
    for config, n, tail in (("c5", 20000, False), ("c4", 30000, True), ("c2", 300, False), ("c3", 500, False),
                            ("c1", 777, True)):
        store, _ = synth.make(config, n=n)
        lens = store.lengths
        pre = np.concatenate([[0], np.cumsum(lens)])
        for nshards in (1, 2, 8):
            plan = plan_strips(store, nshards)
            rows = sorted((a, a + b) for a, b, _ in plan)
            assert rows[0][0] == 0 and rows[-1][1] == store.num
            assert all(rows[k][1] == rows[k + 1][0] for k in range(len(rows) - 1))
            assert all(a % 32 == 0 for a, _, _ in plan)  # a 32-slot query group never straddles a strip
            if n < 10000:
                continue
            load = np.zeros(nshards)
            for a, b, o in plan:
                load[o] += float((lens[a:a + b] * pre[a:a + b]).sum())
            assert load.min() > 0 and load.max() / load.mean() < 1.01
            for sh in range(nshards):
                mine = [float((lens[a:a + b] * pre[a:a + b]).sum()) for a, b, o in plan if o == sh]
                if tail:
                    assert len(mine) >= 5 and min(mine) < 0.1 * sum(mine)
                else:
                    assert len(mine) == (4 if nshards == 1 else 2) and min(mine) > 0.2 * sum(mine)
    store, _ = synth.make("c5", n=20000)
    assert plan_strips(store, 4) == plan_strips(store, 4)  # deterministic on every rank
    uniform = plan_strips(store, 1, panel=1024)
    assert [b for _, b, _ in uniform[:-1]] == [1024] * (len(uniform) - 1)


def test_filter_resolve_is_the_sequential_recurrence(oracle_mod):
    """sab_filter_resolve (host only): fed with every pair that passes the threshold -- computed
    here with numpy in the reference's float32 arithmetic -- it must reproduce filter.c:35-55."""
    import numpy as np
    from sequencealigner_b200 import synth
    from sequencealigner_b200.aligner import filter_resolve
    store, _ = synth.make("c4", n=700)
    seqs = [np.frombuffer(store.sequence(k), dtype=np.uint8) for k in range(store.num)]
    for thr in (0.9, 0.4, 0.05):
        pairs = []
        for j in range(1, store.num):
            for i in range(j):
                ml = min(len(seqs[i]), len(seqs[j]))
                m = int((seqs[i][:ml] == seqs[j][:ml]).sum())
                if np.float32(m) / np.float32(ml) >= np.float32(thr):
                    pairs.append((j << 32) | i)
        rng = np.random.default_rng(1)
        pairs = np.array(pairs, dtype=np.uint64)
        rng.shuffle(pairs)  # phase A delivers them in no particular order
        got = filter_resolve(store.num, pairs)
        assert np.array_equal(got, oracle_mod.filter_lost(store, thr)), thr
    assert not filter_resolve(5, np.zeros(0, np.uint64)).any()


@pytest.mark.parametrize("method,packed,width,value", [
    (1, 1, 32, 0xfff7fff7),  # GA int16x2: lane(o - e) = -9 in both lanes (-s 10 -e 1)
    (2, 1, 30, 0xffffffff),  # SW int16x2: lane(e) = -1
    (2, 0, 16, 0xfffffffe),  # SW int32: e = -2
    (1, 0, 32, 0xffffff9c),  # GA int32: o - e = -100
    (0, 1, 24, 0),           # NW has no gap constant in its add-max ops
])
def test_kernel_images_are_specialised_by_immediate_only(tmp_path, method, packed, width, value):
    """The strip kernels ship as cubins with a placeholder immediate (csrc/sab_module.cu).  Writing a
    job's gap constant must change exactly the immediate operands of VIADDMNMX instructions of ONE
    kernel and nothing else -- checked on the disassembly, which needs no GPU."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    lib = capi.load()
    size = C.c_size_t(0)
    name = C.create_string_buffer(200)
    assert lib.sab_kernel_image(method, packed, width, value, None, 0, C.byref(size), name, 200) == 0
    assert b"sab_wave_kernel" in name.value and size.value > 100000
    placeholder = 0x5ab75ab7

    def sass(val):
        buf = (C.c_ubyte * size.value)()
        sites = lib.sab_kernel_image(method, packed, width, val, buf, size.value, C.byref(size), name, 200)
        path = tmp_path / f"img_{val:x}.cubin"
        path.write_bytes(bytes(buf))
        txt = subprocess.run(["cuobjdump", "-sass", "-fun", name.value.decode(), str(path)],
                             capture_output=True, text=True, check=True).stdout
        ins = [l.split("*/", 1)[1].split(";")[0].strip() for l in txt.splitlines()
               if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
        return sites, ins, bytes(buf)

    sites0, ins0, raw0 = sass(placeholder)  # identity patch: the image as built
    sites1, ins1, raw1 = sass(value)
    assert sites0 == sites1 and len(ins0) == len(ins1) > 1000
    if method == 0:
        assert sites1 == 0 and raw0 == raw1
        return
    # GA packed: 2 add-max ops per cell pair in the 8-, 4-, 2- and 1-row blocks; SW: 8-row block only
    # (SW int32 keeps the vertical gap in __viaddmax_s32_relu with the raw extend: one site per cell)
    rows = 15 if method == 1 else 8
    assert sites1 == (1 if (method == 2 and not packed) else 2) * width * rows
    diff = [(a, b) for a, b in zip(ins0, ins1) if a != b]
    assert len(diff) == sites1
    for a, b in diff:
        assert a.startswith("VIADDMNMX") and f"0x{placeholder:x}" in a
        assert b == a.replace(f"0x{placeholder:x}", f"0x{value:x}")
    assert not any(f"0x{placeholder:x}" in b for b in ins1)
    # every other kernel of the image still carries its placeholder (they are specialised separately)
    assert sum(x != y for x, y in zip(raw0, raw1)) <= 4 * sites1
    assert lib.sab_kernel_image(method, 0, 24, value, None, 0, C.byref(size), name, 200) < 0


@pytest.mark.parametrize("config,n,bits", [("c3", 700, 2), ("c2", 300, 5), ("c5", 500, 5), ("c4", 900, 5)])
def test_upload_packing_roundtrip(config, n, bits):
    """sab_upload's device layouts read back on the host with the kernels' addressing: the subject bit
    stream (2 bits per letter for ACGT reads, 5 for protein; a 5-bit letter may straddle two words) and
    the lane-interleaved query words, for every letter of every sequence (no device needed)."""
    store, c = synth.make(config, n=n)
    cfg = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"), gap_open=c.get("gap_open"),
                      gap_ext=c.get("gap_ext"))
    lib = capi.load()
    r = cfg.resolve()
    sc = capi.SabConfig()
    sc.method = int(r["method"])
    sc.gap_pen, sc.gap_open, sc.gap_ext = int(r["gap_pen"]), int(r["gap_open"]), int(r["gap_ext"])
    sc.seq_lut[:] = [int(v) for v in np.asarray(r["seq_lut"]).reshape(128)]
    sc.sub_mat[:] = [int(v) for v in np.asarray(r["sub_mat"]).reshape(576)]
    got = C.c_int32(0)
    meta = store.meta.ctypes.data_as(C.POINTER(capi.SabMeta))
    bad = lib.sab_selftest_packing(C.byref(sc), store.seqs.ctypes.data, meta, store.num, C.byref(got))
    assert bad == 0 and got.value == bits
