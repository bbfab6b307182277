"""Minimal HDF5 reader used to check sab_h5_write (tests only).

Independent of the writer: it walks the file the way the HDF5 File Format
Specification describes it -- superblock (version 0) -> root symbol-table entry ->
group B-tree (v1) -> symbol table nodes -> local heap names -> version-1 object
headers -> dataspace / datatype / layout / filter-pipeline messages -> contiguous
data, chunk B-tree (v1) + deflate, global heap collections for variable-length
strings -- and validates signatures, sizes, alignment and key ordering on the way.
No libhdf5 / h5py exists in this image; this is the reader-side structural test
SURVEY.md section 8(f1) asks for.

The reader itself is pinned against a file that libhdf5 DID write: a MATLAB v7.3 .mat file
(HDF5 behind a 512-byte user block) that ships with SciPy's test data and is committed as
tests/golden/ext_libhdf5_matlab73_testdouble.mat.  `File(path, strict=False)` walks it through
the same code -- superblock v0 with a base address, root symbol-table entry, group B-tree,
symbol-table node, local heap, version-1 object headers (with messages the writer never emits
skipped), dataspace, datatype, contiguous layout -- see tests/test_h5.py.  What that file does
not contain (chunk B-tree, filter pipeline, global-heap strings) stays checked against the
specification only.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


class H5Error(AssertionError):
    pass


def need(cond, msg):
    if not cond:
        raise H5Error(msg)


class File:
    def __init__(self, path, strict=True):
        """strict: the file must look exactly like sab_h5_write's output (no user block, only the
        messages it emits).  strict=False also accepts what other libhdf5-era writers produce:
        a user block (superblock at 512, 1024, ...; addresses relative to the base address),
        unknown header messages, header continuation blocks, layout message versions 1-2."""
        raw = open(path, "rb").read()
        self.strict = strict
        off = 0
        if not strict:
            while raw[off:off + 8] != SIG and off < len(raw):
                off = 512 if off == 0 else off * 2  # the specification's search sequence
        need(raw[off:off + 8] == SIG, "bad signature")
        self.b = b = raw[off:]  # every address in the file is relative to the base address
        need(b[8] == 0 and b[9] == 0 and b[10] == 0 and b[12] == 0, "superblock versions")
        need(b[13] == 8 and b[14] == 8, "offset/length sizes")
        self.leaf_k, self.int_k = struct.unpack_from("<HH", b, 16)
        base, free, eof, drv = struct.unpack_from("<QQQQ", b, 24)
        need(base == off and free == UNDEF and drv == UNDEF, "superblock addresses")
        need(eof == len(raw), f"end-of-file address {eof} != file size {len(raw)}")
        name_off, self.root_hdr, cache, _, self.root_btree, self.root_heap = \
            struct.unpack_from("<QQIIQQ", b, 56)
        need(cache == 1 and name_off == 0, "root entry")

    # ---- groups ---------------------------------------------------------------------------
    def heap_names(self, addr):
        b = self.b
        need(b[addr:addr + 4] == b"HEAP" and b[addr + 4] == 0, "local heap signature")
        size, free, data = struct.unpack_from("<QQQ", b, addr + 8)
        need(free == 1 or free < size, "local heap free list")  # 1 = H5HL_FREE_NULL
        need(b[data] == 0, "heap offset 0 must be the empty string")
        return lambda off: b[data + off:b.index(b"\0", data + off)].decode()

    def links(self):
        """{name: object header address} of the root group."""
        msgs = self.messages(self.root_hdr)
        st = [d for t, d in msgs if t == 0x11]
        need(len(st) == 1, "root group has no symbol table message")
        btree, heap = struct.unpack_from("<QQ", st[0], 0)
        need((btree, heap) == (self.root_btree, self.root_heap), "root scratch-pad mismatch")
        name = self.heap_names(heap)
        out = {}

        def walk(addr):
            b = self.b
            need(b[addr:addr + 4] == b"TREE" and b[addr + 4] == 0, "group B-tree node")
            level, used = b[addr + 5], struct.unpack_from("<H", b, addr + 6)[0]
            need(used <= 2 * self.int_k, "too many entries")
            p = addr + 24
            keys = []
            for i in range(used):
                key, child = struct.unpack_from("<QQ", b, p)
                keys.append(name(key))
                p += 16
                if level:
                    walk(child)
                else:
                    snod(child)
            keys.append(name(struct.unpack_from("<Q", b, p)[0]))
            need(keys == sorted(keys), "group keys not ordered")
            return keys

        def snod(addr):
            b = self.b
            need(b[addr:addr + 4] == b"SNOD" and b[addr + 4] == 1, "symbol table node")
            n = struct.unpack_from("<H", b, addr + 6)[0]
            need(n <= 2 * self.leaf_k, "too many symbols")
            names = []
            for i in range(n):
                noff, hdr, cache = struct.unpack_from("<QQI", b, addr + 8 + 40 * i)
                out[name(noff)] = hdr
                names.append(name(noff))
            need(names == sorted(names), "symbols not sorted")

        keys = walk(btree)
        need(max(out) <= keys[-1], "last key below the largest name")
        return out

    # ---- object headers -----------------------------------------------------------------------
    def messages(self, addr):
        b = self.b
        need(addr % 8 == 0, "object header alignment")
        ver, _, n, ref, size = struct.unpack_from("<BBHII", b, addr)
        need(ver == 1 and ref >= 1, "object header version")
        p, end, out = addr + 16, addr + 16 + size, []
        blocks = []
        while len(out) < n:
            if p >= end:  # next continuation block (message 0x10), none in sab_h5_write's files
                need(blocks and not self.strict, "message count / header size mismatch")
                p, ln = blocks.pop(0)
                end = p + ln
            t, sz, flags = struct.unpack_from("<HHB", b, p)
            need(sz % 8 == 0, "message size not a multiple of 8")
            out.append((t, b[p + 8:p + 8 + sz]))
            if t == 0x10:
                blocks.append(struct.unpack_from("<QQ", b, p + 8))
            p += 8 + sz
        need(p == end or not self.strict, "message count / header size mismatch")
        return out

    def dataset(self, addr):
        info = {}
        for t, d in self.messages(addr):
            if t == 0x1:
                need(d[0] == 1, "dataspace version")
                rank = d[1]
                info["shape"] = struct.unpack_from("<" + "Q" * rank, d, 8)
            elif t == 0x3:
                cls, ver = d[0] & 15, d[0] >> 4
                need(ver == 1, "datatype version")
                size = struct.unpack_from("<I", d, 4)[0]
                if cls == 0:
                    need(d[1] & 1 == 0 and d[1] & 8 and size == 4, "expected int32 LE signed")
                    need(struct.unpack_from("<HH", d, 8) == (0, 32), "bit offset/precision")
                    info["dtype"] = "<i4"
                elif cls == 9:
                    need(d[1] & 15 == 1 and size == 16, "expected a variable-length string")
                    need(d[8] == 0x13 and struct.unpack_from("<I", d, 12)[0] == 1, "base type")
                    info["dtype"] = "vlstr"
                elif cls == 1 and not self.strict:  # IEEE float, little endian
                    need(d[1] & 1 == 0 and size in (4, 8), "expected a little-endian float")
                    info["dtype"] = "<f%d" % size
                else:
                    raise H5Error(f"datatype class {cls}")
            elif t == 0x8 and d[0] in (1, 2) and not self.strict:
                # layout message versions 1-2: dimensionality, class, 5 reserved, address, dims + element size
                need(d[2] == 1, "only contiguous storage in the old layout message")
                nd = d[1]
                dims = struct.unpack_from("<" + "I" * nd, d, 16)
                info["layout"] = ("contiguous", struct.unpack_from("<Q", d, 8)[0], int(np.prod(dims)))
            elif t == 0x8:
                need(d[0] == 3, "layout version")
                if d[1] == 1:
                    info["layout"] = ("contiguous",) + struct.unpack_from("<QQ", d, 2)
                elif d[1] == 2:
                    nd = d[2]
                    btree = struct.unpack_from("<Q", d, 3)[0]
                    dims = struct.unpack_from("<" + "I" * nd, d, 11)
                    info["layout"] = ("chunked", btree, dims)
                else:
                    raise H5Error("layout class")
            elif t == 0xB:
                need(d[0] == 1 and d[1] == 1, "filter pipeline")
                fid, nlen, flags, ncd = struct.unpack_from("<HHHH", d, 8)
                need(fid == 1 and nlen == 0 and ncd == 1, "expected deflate")
                info["deflate"] = struct.unpack_from("<I", d, 16)[0]
            elif t == 0x5:
                need((d[0] == 2 and d[3] == 0) or not self.strict, "fill value message")
        need({"shape", "dtype", "layout"} <= set(info), "incomplete dataset header")
        return info

    # ---- data -------------------------------------------------------------------------------------
    def read(self, addr):
        info = self.dataset(addr)
        if info["dtype"] == "vlstr":
            kind, data, size = info["layout"]
            n = info["shape"][0]
            need(kind == "contiguous" and size == 16 * n, "vlstr layout")
            return [self.heap_object(*struct.unpack_from("<IQI", self.b, data + 16 * k))
                    for k in range(n)], info
        shape = info["shape"]
        if info["layout"][0] == "contiguous":
            _, data, size = info["layout"]
            item = np.dtype(info["dtype"]).itemsize
            need(size == item * shape[0] * shape[1], "contiguous size")
            return np.frombuffer(self.b, info["dtype"], shape[0] * shape[1], data).reshape(shape), info
        _, btree, cd = info["layout"]
        need(len(cd) == 3 and cd[2] == 4 and cd[0] == cd[1], "chunk dims")
        out = np.full(shape, -7777, dtype=np.int32)
        seen = []
        self.chunks(btree, None, seen)
        need(seen == sorted(seen), "chunks not in key order")
        need(len(seen) == -(-shape[0] // cd[0]) * -(-shape[1] // cd[1]), "chunk count")
        for (r, c), nbytes, a in self._chunk_list:
            raw = self.b[a:a + nbytes]
            if "deflate" in info:
                raw = zlib.decompress(raw)
            need(len(raw) == cd[0] * cd[1] * 4, "chunk size")
            blk = np.frombuffer(raw, "<i4").reshape(cd[0], cd[1])
            h, w = min(cd[0], shape[0] - r), min(cd[1], shape[1] - c)
            need(not blk[h:, :].any() and not blk[:, w:].any(), "edge padding not zero")
            out[r:r + h, c:c + w] = blk[:h, :w]
        info["nodes"] = self._nodes
        return out, info

    def chunks(self, addr, level_expected, seen):
        if level_expected is None:
            self._chunk_list, self._nodes = [], 0
        b = self.b
        need(addr % 8 == 0 and b[addr:addr + 4] == b"TREE" and b[addr + 4] == 1, "chunk B-tree node")
        self._nodes += 1
        level, used = b[addr + 5], struct.unpack_from("<H", b, addr + 6)[0]
        need(level_expected is None or level == level_expected, "node level")
        need(1 <= used <= 64, "entries used")
        p, keys = addr + 24, []
        for i in range(used + 1):
            nbytes, mask, r, c, z = struct.unpack_from("<IIQQQ", b, p)
            need(mask == 0 and z == 0, "chunk key")
            keys.append((r, c))
            p += 32
            if i < used:
                child = struct.unpack_from("<Q", b, p)[0]
                p += 8
                if level:
                    self.chunks(child, level - 1, seen)
                else:
                    seen.append((r, c))
                    self._chunk_list.append(((r, c), nbytes, child))
        need(keys == sorted(keys) and len(set(keys)) == len(keys), "chunk keys not increasing")
        return keys

    def heap_object(self, length, col, idx):
        b = self.b
        need(b[col:col + 4] == b"GCOL" and b[col + 4] == 1, "global heap collection")
        size = struct.unpack_from("<Q", b, col + 8)[0]
        need(size >= 4096 and size % 8 == 0, "collection size")
        p, end = col + 16, col + size
        while p + 16 <= end:
            oid, ref, _, osz = struct.unpack_from("<HHIQ", b, p)
            if oid == 0:
                need(p + osz == end, "free-space object must cover the rest of the collection")
                break
            if oid == idx:
                need(osz == length, "string length mismatch")
                return b[p + 16:p + 16 + osz]
            p += 16 + ((osz + 7) & ~7)
        raise H5Error(f"heap object {idx} not found")


def read_similarity_file(path):
    """-> (sequences: list[bytes], matrix: int32[dim,dim], info)"""
    f = File(path)
    links = f.links()
    need(sorted(links) == ["sequences", "similarity_matrix"], f"root group holds {sorted(links)}")
    seqs, si = f.read(links["sequences"])
    mat, mi = f.read(links["similarity_matrix"])
    need(si["shape"] == (len(seqs),) and mi["shape"] == (len(seqs), len(seqs)), "shapes")
    return seqs, mat, mi
