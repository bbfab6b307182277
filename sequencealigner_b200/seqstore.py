"""Host-side sequence store: the layout the reference hands to its backend.

Mirrors `struct input {uchar *seqs; struct meta *meta; s32 max; s32 num;}`
(reference: src/io/input.h:6-11, src/bio/align.h:6-9): uppercase ASCII letters,
every sequence NUL-terminated, back to back in one buffer; meta[k] = (off, len).
Parsing follows the reference's FASTA reader (src/io/source/fasta.c:16-85):
letters are upper-cased, blanks and line ends skipped, every letter must be in
the selected matrix's alphabet, empty sequences are errors.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

META_DTYPE = np.dtype([("off", "<i4"), ("len", "<i4")])


@dataclass
class SeqStore:
    seqs: np.ndarray  # uint8, NUL-separated letters
    meta: np.ndarray  # META_DTYPE[num]

    @property
    def num(self) -> int:
        return int(self.meta.shape[0])

    @property
    def max(self) -> int:
        return int(self.meta["len"].max()) if self.num else 0

    @property
    def lengths(self) -> np.ndarray:
        return self.meta["len"].astype(np.int64)

    def pairs(self) -> int:
        n = self.num
        return n * (n - 1) // 2  # alignments(n), src/util/macros.h:13

    def cells(self) -> int:
        """sum_{i<j} len_i*len_j, exactly."""
        l = self.lengths
        s = int(l.sum())
        return (s * s - int((l * l).sum())) // 2

    def sequence(self, k: int) -> bytes:
        o, l = int(self.meta["off"][k]), int(self.meta["len"][k])
        return self.seqs[o:o + l].tobytes()

    def subset(self, keep: np.ndarray) -> "SeqStore":
        """Order-preserving compaction, as filter() does (src/bio/filter.c:64-77)."""
        return SeqStore.from_sequences([self.sequence(int(k)) for k in np.flatnonzero(keep)])

    def head(self, n: int) -> "SeqStore":
        n = min(n, self.num)
        end = int(self.meta["off"][n - 1] + self.meta["len"][n - 1] + 1)
        return SeqStore(self.seqs[:end].copy(), self.meta[:n].copy())

    @staticmethod
    def from_sequences(seqs) -> "SeqStore":
        seqs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
        lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=len(seqs))
        if len(seqs) and lens.min() < 1:
            raise ValueError("empty sequence (reference: fasta.c:64-67)")
        offs = np.zeros(len(seqs), dtype=np.int64)
        if len(seqs):
            offs[1:] = np.cumsum(lens[:-1] + 1)
        total = int((lens + 1).sum())
        if total > 2**31 - 1:
            raise ValueError("length overflow (reference: fasta.c:72-75)")
        buf = np.frombuffer(b"\0".join(seqs) + b"\0", dtype=np.uint8).copy() if seqs else \
            np.zeros(0, np.uint8)
        meta = np.zeros(len(seqs), dtype=META_DTYPE)
        meta["off"] = offs
        meta["len"] = lens
        return SeqStore(buf, meta)

    @staticmethod
    def from_codes(codes: np.ndarray, lens: np.ndarray, alphabet: str) -> "SeqStore":
        """Vectorised constructor for synthetic sets: `codes` are indices into
        `alphabet`, concatenated without separators."""
        lens = np.asarray(lens, dtype=np.int64)
        n = lens.shape[0]
        offs = np.zeros(n, dtype=np.int64)
        offs[1:] = np.cumsum(lens[:-1] + 1)
        total = int((lens + 1).sum())
        buf = np.zeros(total, dtype=np.uint8)
        letters = np.frombuffer(alphabet.encode(), dtype=np.uint8)[codes]
        # position of letter t of sequence k: offs[k] + t
        starts = np.zeros(n, dtype=np.int64)
        starts[1:] = np.cumsum(lens[:-1])
        idx = np.arange(int(lens.sum()), dtype=np.int64)
        seq_id = np.repeat(np.arange(n), lens)
        buf[idx + (offs - starts)[seq_id]] = letters
        meta = np.zeros(n, dtype=META_DTYPE)
        meta["off"] = offs
        meta["len"] = lens
        return SeqStore(buf, meta)

    def to_fasta(self, path) -> None:
        with open(path, "wb") as f:
            for k in range(self.num):
                f.write(b">s%d\n" % k)
                f.write(self.sequence(k))
                f.write(b"\n")


def read_fasta(path, seq_lut=None) -> SeqStore:
    """fasta.c:16-85 semantics.  `seq_lut` (int32[128]) validates the alphabet."""
    data = open(path, "rb").read()
    if not data.startswith(b">"):
        raise ValueError("Data before first header")
    out = []
    for rec in data.split(b">")[1:]:
        nl = rec.find(b"\n")
        if nl < 0:
            raise ValueError("Last header has no data")
        body = rec[nl + 1:].upper().translate(None, b"\r\n ")
        if not body:
            raise ValueError(f"Sequence #{len(out) + 1} is empty")
        arr = np.frombuffer(body, dtype=np.uint8)
        if arr.max() > 127 or arr.min() == 0:
            raise ValueError(f"Sequence #{len(out) + 1} is corrupted")
        if seq_lut is not None and (np.asarray(seq_lut)[arr] < 0).any():
            raise ValueError(f"Sequence #{len(out) + 1} is invalid")
        out.append(body)
    return SeqStore.from_sequences(out)
