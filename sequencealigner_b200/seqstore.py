"""Host-side sequence store: the layout the reference hands to its backend.

Mirrors `struct input {uchar *seqs; struct meta *meta; s32 max; s32 num;}`
(reference: src/io/input.h:6-11, src/bio/align.h:6-9): uppercase ASCII letters,
every sequence NUL-terminated, back to back in one buffer; meta[k] = (off, len).
Files are parsed by the library (sab_parse_fasta / sab_parse_dsv, csrc/sab_source.cpp), which
restates the reference's readers (src/io/source/fasta.c:16-85, src/io/source/dsv.c:80-230):
letters are upper-cased, blanks and line ends skipped, every letter must be in the selected
matrix's alphabet, empty sequences are errors -- with the reference's own message.
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


class ParseError(ValueError):
    """The input text was rejected; the message is the reference's own (perr line)."""


def parse_source(data: bytes, ext: str, seq_lut, gap_pen: int = 0, column: int | None = None,
                 header: bool = True) -> SeqStore:
    """The reference's input_load (src/io/input.c:28-81) on an in-memory file: picks the parser by
    extension, runs it (libseqalign_b200.so: sab_parse_fasta / sab_parse_dsv, which restate
    src/io/source/fasta.c:16-85 and src/io/source/dsv.c:80-230) and builds the meta table.

    `seq_lut` = SEQ_LUT of the selected matrix, `gap_pen` <= 0 = GAP_PEN (bounds the lengths).
    DSV: the sequence column is found by its header name (sequence, seq, protein, ...); where the
    reference would ASK (no such name, dsv.c:146-160) the caller answers through `column`
    (zero-based) and `header` (False = the first row is data, the "No header line" answer)."""
    import ctypes as C

    from . import capi
    lib = capi.load()
    lut = (C.c_int32 * 128)(*[int(v) for v in np.asarray(seq_lut).reshape(128)])
    delim = C.c_uint8(0)
    kind = lib.sab_source_format(ext.encode(), C.byref(delim))
    if kind == capi.SAB_SRC_NONE:
        raise ParseError(f"Unsupported file format: .{ext}")
    buf = np.zeros(len(data) + 1, dtype=np.uint8)  # + 1: see the header's note on headerless DSV
    buf[:len(data)] = np.frombuffer(data, dtype=np.uint8)
    num, longest = C.c_int32(0), C.c_int32(0)
    err = C.create_string_buffer(256)
    if kind == capi.SAB_SRC_FASTA:
        rc = lib.sab_parse_fasta(buf.ctypes.data, len(data), lut, gap_pen, C.byref(num),
                                 C.byref(longest), err, len(err))
    else:
        cols, found = C.c_int32(0), C.c_int32(-1)
        rc = lib.sab_dsv_header(buf.ctypes.data, len(data), delim, C.byref(cols), C.byref(found),
                                None, 0, err, len(err))
        if rc == capi.SAB_OK:
            if column is None:
                if found.value < 0:
                    names = C.create_string_buffer(len(data) + 1)  # the first row's field texts
                    lib.sab_dsv_header(buf.ctypes.data, len(data), delim, C.byref(cols), C.byref(found),
                                       names, len(names), err, len(err))
                    shown = names.raw.split(b"\0")[:cols.value]
                    raise ParseError("no column is named like a sequence column; pass column= "
                                     f"(and header=False if the first row is data): {shown}")
                column = found.value
            if not 0 <= column < cols.value:
                raise ParseError(f"column {column} out of range (the first row has {cols.value})")
            rc = lib.sab_parse_dsv(buf.ctypes.data, len(data), delim, column, int(not header), lut,
                                   gap_pen, C.byref(num), C.byref(longest), err, len(err))
    if rc == capi.SAB_ERR_PARSE:
        raise ParseError(err.value.decode())
    if rc != capi.SAB_OK:
        raise capi.SabError(rc, "input rejected (too large, or gap_pen > 0)")
    if num.value < 2:  # SEQ_N_MIN, src/io/input.c:64-67
        raise ParseError(f"Not enough sequences: {num.value} (min: 2)")
    meta = np.zeros(num.value, dtype=META_DTYPE)
    size = lib.sab_build_meta(buf.ctypes.data, num.value, meta.ctypes.data_as(C.POINTER(capi.SabMeta)))
    store = SeqStore(buf[:size].copy(), meta)
    assert store.max == longest.value
    return store


def read_source(path, seq_lut, gap_pen: int = 0, column: int | None = None,
                header: bool = True) -> SeqStore:
    """input_load for a file on disk: the extension after the last dot selects the parser
    (src/io/input.c:31-36)."""
    from pathlib import Path
    name = Path(path).name
    dot = name.rfind(".")
    if dot <= 0:
        raise ParseError(f"File extension not found: {name}")
    return parse_source(Path(path).read_bytes(), name[dot + 1:], seq_lut, gap_pen, column, header)


def read_fasta(path, seq_lut) -> SeqStore:
    """fasta.c:16-85 on a file of any name."""
    from pathlib import Path
    return parse_source(Path(path).read_bytes(), "fasta", seq_lut)
