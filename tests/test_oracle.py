"""CPU-only: pins the oracle restatement (oracle/seqalign_oracle.c) against
 (1) the committed golden fixtures, which are outputs of the UNMODIFIED reference
     CPU build (tools/make_golden.py), and
 (2) the reference binary itself when it is present (build container only)."""
import numpy as np
import pytest

from conftest import GOLDEN, cli_to_config, golden_cases, load_golden, parse_fasta_bytes
from sequencealigner_b200 import SeqStore, synth


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_matches_golden(oracle_mod, name):
    g = load_golden(name)
    cfg = cli_to_config(g["args"])
    store = SeqStore.from_sequences(g["seqs"])
    if cfg.filter > 0:
        lost = oracle_mod.filter_lost(store, cfg.filter)
        store = store.subset(~lost)
    assert [store.sequence(k) for k in range(store.num)] == g["kept"]
    mat, cells = oracle_mod.align_all(cfg.resolve(), store, triangular=g["triangular"])
    assert cells == store.cells()
    assert np.array_equal(mat, g["matrix"])


def test_oracle_matches_golden_all_tables(oracle_mod):
    z = np.load(GOLDEN / "all_tables.npz")
    stores = {k: SeqStore.from_sequences(parse_fasta_bytes(z[k])) for k in ("protein", "nucleo")}
    from sequencealigner_b200 import AlignConfig
    n = 0
    for key in z.files:
        if "/" not in key:
            continue
        name, method = key.split("/")
        cfg = AlignConfig(method=method, matrix=name, gap_pen=3) if method == "nw" else \
            AlignConfig(method=method, matrix=name, gap_open=9, gap_ext=2)
        store = stores["nucleo" if name in ("dnafull", "nuc44") else "protein"]
        mat, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=False)
        assert np.array_equal(mat, z[key]), key
        n += 1
    assert n == 67 * 3


def test_oracle_properties(oracle_mod):
    from sequencealigner_b200 import AlignConfig
    store, _ = synth.make("c1", n=40)
    r = AlignConfig("sw", "blosum62", gap_open=10, gap_ext=1).resolve()
    full, _ = oracle_mod.align_all(r, store, triangular=False)
    tri, _ = oracle_mod.align_all(r, store, triangular=True)
    assert (full == full.T).all() and (np.diag(full) == 0).all() and (full >= 0).all()
    n = store.num
    for j in range(1, n):
        assert np.array_equal(tri[j * (j - 1) // 2: j * (j - 1) // 2 + j], full[j, :j])
    # a sequence against itself under NW scores the sum of its diagonal substitution scores
    r = AlignConfig("nw", "blosum62", gap_pen=4).resolve()
    s = store.sequence(3)
    want = int(sum(r["sub_mat"][r["seq_lut"][c], r["seq_lut"][c]] for c in s))
    assert oracle_mod.pair(r, s, s) == want


@pytest.mark.skipif(not __import__("pathlib").Path("/root/reference/src").is_dir(),
                    reason="reference sources only exist in the build container")
@pytest.mark.parametrize("method,kw", [("nw", dict(gap_pen=4)), ("ga", dict(gap_open=10, gap_ext=1)),
                                       ("sw", dict(gap_open=10, gap_ext=1))])
def test_oracle_matches_reference_binary(oracle_mod, tmp_path, method, kw):
    from sequencealigner_b200 import AlignConfig
    assert oracle_mod.ref_available()
    store, _ = synth.make("c2", n=24, seed=7)
    fa = tmp_path / "x.fasta"
    store.to_fasta(fa)
    ref = oracle_mod.run_reference(fa, oracle_mod.ref_args(method, "blosum62", **kw), threads=2)
    mat, _ = oracle_mod.align_all(AlignConfig(method, "blosum62", **kw).resolve(), store,
                                  triangular=ref["triangular"])
    assert np.array_equal(mat, ref["matrix"])


@pytest.mark.skipif(not __import__("pathlib").Path("/root/reference/src").is_dir(),
                    reason="reference sources only exist in the build container")
def test_oracle_filter_matches_reference_binary(oracle_mod, tmp_path):
    store, cfg = synth.make("c4", n=300)
    fa = tmp_path / "f.fasta"
    store.to_fasta(fa)
    ref = oracle_mod.run_reference(fa, oracle_mod.ref_args("ga", "blosum62", gap_open=10, gap_ext=1,
                                                           filter=0.9), threads=1)
    lost = oracle_mod.filter_lost(store, 0.9)
    assert 0 < lost.sum() < store.num
    kept = store.subset(~lost)
    assert [kept.sequence(k) for k in range(kept.num)] == ref["seqs"]


@pytest.mark.parametrize("name,npairs", [("c5_full_200k_8gpu_samples", 1500), ("c5ovf_20k_8gpu_samples", 1200)])
def test_full_size_gpu_run_samples_equal_the_oracle(oracle_mod, name, npairs):
    """Scores the B200 path wrote during the FULL-SIZE runs of round 2 (C5: 200 000 sequences on 8 GPUs
    into a file-backed sink; C5-ovf: 20 000 sequences with planted poly-W under pam500), sampled on the
    GPU box by tools/full_run.py and committed as a fixture: a subset is re-scored here by the CPU oracle
    (all 10^6 sampled pairs + 18 whole rows of either run were checked once by tools/verify_samples.py:
    profiles/r2_verify_samples.json)."""
    import numpy as np
    from pathlib import Path
    from sequencealigner_b200 import AlignConfig, synth
    z = np.load(Path(__file__).parent / "golden" / "gpu_samples" / f"{name}.npz")
    n_in = int(z["sequences_in"])
    if str(z["config"]).endswith("-ovf"):
        store, c = synth.make_overflow(n=n_in, planted=64)
    else:
        store, c = synth.make(str(z["config"]), n=n_in)
    assert store.num == int(z["n"])
    r = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"), gap_open=c.get("gap_open"),
                    gap_ext=c.get("gap_ext")).resolve()
    i, j, score = z["i"], z["j"], z["score"]
    pick = np.argsort(-score.astype(np.int64), kind="stable")[:200]  # the highest scores (int16 overflows) ...
    pick = np.unique(np.concatenate([pick, np.arange(npairs)]))       # ... and a prefix of the uniform sample
    want = oracle_mod.pairs(r, store, i[pick], j[pick])
    assert np.array_equal(want, score[pick])
    if "ovf" in name:
        assert score.max() == 34000  # poly-W x poly-W under pam500: beyond int16, re-scored in int32
