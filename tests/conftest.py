import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def parse_fasta_bytes(buf) -> list[bytes]:
    data = bytes(np.asarray(buf, dtype=np.uint8).tobytes())
    return [rec.split(b"\n", 1)[1].replace(b"\n", b"") for rec in data.split(b">")[1:]]


def cli_to_config(args):
    """['-m','blosum62','-a','ga','-s','10','-e','1','-f','0.9'] -> AlignConfig"""
    from sequencealigner_b200 import AlignConfig
    a = [str(x) for x in args]
    kv = dict(zip(a[0::2], a[1::2]))
    return AlignConfig(method=kv["-a"], matrix=kv["-m"],
                       gap_pen=int(kv["-p"]) if "-p" in kv else None,
                       gap_open=int(kv["-s"]) if "-s" in kv else None,
                       gap_ext=int(kv["-e"]) if "-e" in kv else None,
                       filter=float(kv.get("-f", 0.0)))


def golden_cases():
    return sorted(p.stem for p in GOLDEN.glob("*.npz") if p.stem != "all_tables")


def load_golden(name):
    z = np.load(GOLDEN / f"{name}.npz", allow_pickle=False)
    return dict(seqs=parse_fasta_bytes(z["fasta"]), args=[str(x) for x in z["args"]],
                matrix=z["matrix"], triangular=bool(z["triangular"]),
                kept=[bytes(x) for x in z["kept"]])


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle
    oracle.build(ref=(Path("/root/reference/src").is_dir()))
    return oracle
