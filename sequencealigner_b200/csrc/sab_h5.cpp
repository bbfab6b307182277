// sab_h5.cpp -- self-contained HDF5 writer for the similarity-matrix file (host only).
//
// Writes the file the reference produces with libhdf5 in flush_hdf5()
// (reference: src/io/format/hdf5.c:14-202): two datasets in the root group,
//
//   /sequences          1-D [dim]      variable-length C strings (H5T_C_S1, H5T_VARIABLE)
//   /similarity_matrix  2-D [dim,dim]  H5T_STD_I32LE; contiguous when dim <= 256, else
//                       chunked c x c (c = largest power of two <= dim, clamped to
//                       [256, 4096] -- what hdf5.c:68-84 evaluates to) with deflate level
//                       `compression` when > 0; a packed-triangular input is expanded to the
//                       full symmetric matrix chunk by chunk (hdf5.c:149-194)
//
// without linking libhdf5 (absent from the build image).  The container structures are
// the original, universally readable ones of the HDF5 File Format Specification:
// superblock version 0, version-1 object headers, symbol-table group (B-tree v1 + local
// heap + SNOD), layout message version 3, B-tree v1 chunk index, filter pipeline
// version 1, global heap collections for the variable-length strings.  The reference asks
// libhdf5 for "latest" structures; the datasets, types, shapes, chunking, filter and values
// a reader sees are the same.
//
// Differences that are deliberate: the diagonal is written as 0 (the reference leaves the
// diagonal cells of rows past its first row block unassigned, hdf5.c:150-162), and chunks
// are deflated by a pool of threads instead of one.
#include "../../include/seqalign_b200.h"

#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <sys/mman.h>
#include <unistd.h>

namespace {

constexpr uint64_t UNDEF = ~0ull;
constexpr int kGroupLeafK = 4, kGroupInternalK = 16, kChunkK = 32; // superblock v0 defaults

struct Buf {
	std::vector<uint8_t> b;
	void u8(unsigned v) { b.push_back((uint8_t)v); }
	void u16(unsigned v) { u8(v & 0xff); u8(v >> 8); }
	void u32(uint32_t v) { u16(v & 0xffff); u16(v >> 16); }
	void u64(uint64_t v) { u32((uint32_t)v); u32((uint32_t)(v >> 32)); }
	void raw(const void *p, size_t n) { b.insert(b.end(), (const uint8_t *)p, (const uint8_t *)p + n); }
	void zero(size_t n) { b.insert(b.end(), n, 0); }
	void pad8() { while (b.size() & 7) b.push_back(0); }
	size_t size() const { return b.size(); }
};

struct File {
	FILE *f = nullptr;
	uint64_t eof = 0;
	bool ok = true;
	void at(uint64_t off, const void *p, size_t n)
	{
		if (!ok || !n)
			return;
		if (fseeko(f, (off_t)off, SEEK_SET) != 0 || fwrite(p, 1, n, f) != n)
			ok = false;
	}
	uint64_t append(const void *p, size_t n)
	{
		const uint64_t off = eof;
		at(off, p, n);
		eof += n;
		return off;
	}
	uint64_t reserve(size_t n) // zero-filled, 8-byte aligned
	{
		eof = (eof + 7) & ~7ull;
		std::vector<uint8_t> z(n, 0);
		return append(z.data(), n);
	}
};

// one version-1 header message: type, flags 0, data padded to a multiple of 8
void message(Buf &h, unsigned type, const Buf &data)
{
	const size_t n = (data.size() + 7) & ~(size_t)7;
	h.u16(type);
	h.u16((unsigned)n);
	h.u8(0);
	h.zero(3);
	h.raw(data.b.data(), data.size());
	h.zero(n - data.size());
}

// version-1 object header around `nmsg` messages
Buf object_header(const Buf &msgs, unsigned nmsg)
{
	Buf h;
	h.u8(1);
	h.u8(0);
	h.u16(nmsg);
	h.u32(1); // reference count
	h.u32((uint32_t)msgs.size());
	h.zero(4); // messages start 8-byte aligned
	h.raw(msgs.b.data(), msgs.size());
	return h;
}

Buf dataspace(const std::vector<uint64_t> &dims)
{
	Buf d;
	d.u8(1); // version
	d.u8((unsigned)dims.size());
	d.u8(0); // no max dims, no permutation
	d.zero(5);
	for (uint64_t v : dims)
		d.u64(v);
	return d;
}

Buf datatype_i32le()
{
	Buf d;
	d.u8(0x10); // version 1, class 0 fixed point
	d.u8(0x08); // little endian, signed
	d.u8(0);
	d.u8(0);
	d.u32(4);
	d.u16(0); // bit offset
	d.u16(32); // precision
	return d;
}

Buf datatype_vlstr()
{
	Buf d;
	d.u8(0x19); // version 1, class 9 variable length
	d.u8(0x01); // type: string; padding: null terminated
	d.u8(0x00); // character set: ASCII
	d.u8(0);
	d.u32(16); // element: length(4) + collection address(8) + object index(4)
	// base type: H5T_C_S1 (one-byte string, null terminated, ASCII)
	d.u8(0x13);
	d.u8(0);
	d.u8(0);
	d.u8(0);
	d.u32(1);
	return d;
}

Buf fill_value(bool chunked)
{
	Buf d;
	d.u8(2); // version
	d.u8(chunked ? 3 : 2); // space allocation: incremental / late
	d.u8(2); // write fill value only if set
	d.u8(0); // undefined (library default 0)
	return d;
}

size_t chunk_dim_for(size_t dim)
{
	if (dim <= 256) // H5_MIN_CHUNK_SIZE, hdf5.c:10-11,69
		return 0;
	size_t c = 64;
	while (c * 2 <= dim)
		c *= 2;
	return std::min<size_t>(std::min<size_t>(std::max<size_t>(c, 256), 4096), dim);
}

struct ChunkRec {
	uint64_t addr;
	uint32_t bytes;
	uint64_t r, c; // element offsets
};

struct Key {
	uint32_t bytes;
	uint64_t r, c;
};

constexpr size_t kChunkNodeBytes = 24 + (2 * kChunkK + 1) * 32 + 2 * kChunkK * 8;

// one B-tree v1 node of the chunk index (type 1), full on-disk size for 2K children
void write_chunk_node(File &F, uint64_t addr, int level, uint64_t left, uint64_t right,
		      const std::vector<Key> &keys, const std::vector<uint64_t> &child)
{
	Buf n;
	n.raw("TREE", 4);
	n.u8(1);
	n.u8((unsigned)level);
	n.u16((unsigned)child.size());
	n.u64(left);
	n.u64(right);
	for (size_t i = 0; i <= child.size(); i++) {
		n.u32(keys[i].bytes);
		n.u32(0); // filter mask: every filter applied
		n.u64(keys[i].r);
		n.u64(keys[i].c);
		n.u64(0);
		if (i < child.size())
			n.u64(child[i]);
	}
	n.zero(kChunkNodeBytes - n.size());
	F.at(addr, n.b.data(), n.size());
}

// Builds the chunk index bottom-up; the nodes of one level are laid out back to back so
// their sibling links are known up front.  Returns the root's address.
uint64_t write_chunk_tree(File &F, const std::vector<ChunkRec> &chunks, uint64_t cdim)
{
	struct Sub {
		uint64_t addr;
		Key first, last;
	};
	std::vector<Sub> cur;
	for (const ChunkRec &c : chunks) // level "-1": the chunks themselves
		cur.push_back({ c.addr, { c.bytes, c.r, c.c }, { 0, c.r, c.c + cdim } });
	const size_t fan = 2 * kChunkK;
	int level = 0;
	do {
		const size_t nn = (cur.size() + fan - 1) / fan;
		F.eof = (F.eof + 7) & ~7ull;
		const uint64_t base = F.eof;
		F.eof += nn * kChunkNodeBytes;
		std::vector<Sub> up;
		for (size_t q = 0; q < nn; q++) {
			const size_t i = q * fan, e = std::min(cur.size(), i + fan);
			std::vector<Key> keys;
			std::vector<uint64_t> child;
			for (size_t k = i; k < e; k++) {
				keys.push_back(cur[k].first);
				child.push_back(cur[k].addr);
			}
			keys.push_back(cur[e - 1].last); // > every chunk below this node
			const uint64_t addr = base + q * kChunkNodeBytes;
			write_chunk_node(F, addr, level, q ? addr - kChunkNodeBytes : UNDEF,
					 q + 1 < nn ? addr + kChunkNodeBytes : UNDEF, keys, child);
			up.push_back({ addr, keys[0], cur[e - 1].last });
		}
		cur.swap(up);
		level++;
	} while (cur.size() > 1);
	return cur[0].addr;
}

inline int32_t cell_at(const int32_t *m, bool tri, uint64_t dim, uint64_t r, uint64_t c)
{
	if (!tri)
		return m[r * dim + c];
	if (r == c)
		return 0;
	const uint64_t hi = std::max(r, c), lo = std::min(r, c);
	return m[hi * (hi - 1) / 2 + lo]; // src/io/output.c:83
}

} // namespace

extern "C" int sab_h5_chunk_dim(size_t dim)
{
	return (int)chunk_dim_for(dim);
}

extern "C" int sab_h5_write(const char *path, const char *const *seqs, size_t dim, const int32_t *matrix,
			    int triangular, unsigned compression, sab_h5_progress_fn progress, void *user)
{
	return sab_h5_write_chunked(path, seqs, dim, matrix, triangular, compression, -1, progress, user);
}

extern "C" int sab_h5_write_chunked(const char *path, const char *const *seqs, size_t dim,
				    const int32_t *matrix, int triangular, unsigned compression,
				    int chunk_dim, sab_h5_progress_fn progress, void *user)
{
	if (!path || !seqs || !matrix || dim < 1 || compression > 9 || chunk_dim > (int)dim)
		return SAB_ERR_ARG;
	File F;
	F.f = fopen(path, "w+b"); // read/write: the uncompressed chunks are written through a shared mapping
	if (!F.f)
		return SAB_ERR_ARG;

	// ---- fixed metadata block: addresses are known before anything is written ----------
	const uint64_t a_super = 0, a_root = 96;
	const size_t cdim = chunk_dim < 0 ? chunk_dim_for(dim) : (size_t)chunk_dim;
	const bool chunked = cdim != 0, deflate = chunked && compression > 0;

	Buf heap_data; // local heap data segment: "", "sequences", "similarity_matrix"
	heap_data.zero(8);
	const uint64_t n_seq = heap_data.size();
	heap_data.raw("sequences", 10);
	heap_data.pad8();
	const uint64_t n_mat = heap_data.size();
	heap_data.raw("similarity_matrix", 18);
	heap_data.pad8();

	const size_t root_hdr_size = 16 + 8 + 16;
	const uint64_t a_heap = a_root + root_hdr_size; // 136
	const uint64_t a_heap_data = a_heap + 32;
	const uint64_t a_btree = a_heap_data + heap_data.size();
	const size_t group_node_size = 24 + (2 * kGroupInternalK + 1) * 8 + 2 * kGroupInternalK * 8;
	const uint64_t a_snod = a_btree + group_node_size;
	const size_t snod_size = 8 + 2 * kGroupLeafK * 40;
	const uint64_t a_dset_seq = a_snod + snod_size;

	// /sequences header (contiguous)
	auto seq_header = [&](uint64_t data_addr) {
		Buf m;
		message(m, 0x0001, dataspace({ (uint64_t)dim }));
		message(m, 0x0003, datatype_vlstr());
		message(m, 0x0005, fill_value(false));
		Buf lay;
		lay.u8(3);
		lay.u8(1); // contiguous
		lay.u64(data_addr);
		lay.u64((uint64_t)dim * 16);
		message(m, 0x0008, lay);
		return object_header(m, 4);
	};
	// /similarity_matrix header
	auto mat_header = [&](uint64_t addr) {
		Buf m;
		unsigned n = 4;
		message(m, 0x0001, dataspace({ (uint64_t)dim, (uint64_t)dim }));
		message(m, 0x0003, datatype_i32le());
		message(m, 0x0005, fill_value(chunked));
		Buf lay;
		lay.u8(3);
		if (chunked) {
			lay.u8(2);
			lay.u8(3); // rank + 1
			lay.u64(addr); // chunk B-tree
			lay.u32((uint32_t)cdim);
			lay.u32((uint32_t)cdim);
			lay.u32(4);
		} else {
			lay.u8(1);
			lay.u64(addr);
			lay.u64((uint64_t)dim * dim * 4);
		}
		message(m, 0x0008, lay);
		if (deflate) {
			Buf fp;
			fp.u8(1);
			fp.u8(1); // one filter
			fp.zero(6);
			fp.u16(1); // H5Z_FILTER_DEFLATE
			fp.u16(0); // no name
			fp.u16(0); // mandatory
			fp.u16(1); // one client value
			fp.u32(compression);
			fp.u32(0); // pad (odd number of values)
			message(m, 0x000B, fp);
			n++;
		}
		return object_header(m, n);
	};
	const size_t seq_hdr_size = seq_header(0).size();
	const uint64_t a_dset_mat = a_dset_seq + seq_hdr_size;
	const size_t mat_hdr_size = mat_header(0).size();
	F.eof = a_dset_mat + mat_hdr_size;
	{ // claim the metadata block
		std::vector<uint8_t> z(F.eof, 0);
		F.at(0, z.data(), z.size());
	}

	// ---- /sequences: global heap collections + the 16-byte references ------------------
	Buf refs;
	{
		size_t k = 0;
		while (k < dim) {
			Buf col; // objects of one collection
			size_t first = k;
			unsigned idx = 1;
			while (k < dim && idx < 60000 && col.size() < (1u << 20)) {
				const size_t len = strlen(seqs[k]);
				col.u16(idx++);
				col.u16(1);
				col.u32(0);
				col.u64(len);
				col.raw(seqs[k], len);
				col.pad8();
				k++;
			}
			size_t total = 16 + col.size();
			size_t want = std::max<size_t>(4096, (total + 16 + 7) & ~(size_t)7);
			Buf g;
			g.raw("GCOL", 4);
			g.u8(1);
			g.zero(3);
			g.u64(want);
			g.raw(col.b.data(), col.size());
			const size_t freesz = want - total; // >= 16 by construction
			g.u16(0);
			g.u16(0);
			g.u32(0);
			g.u64(freesz); // free-space object: size includes its own header
			g.zero(want - g.size());
			F.eof = (F.eof + 7) & ~7ull;
			const uint64_t a_col = F.append(g.b.data(), g.size());
			unsigned id = 1;
			for (size_t q = first; q < k; q++) {
				refs.u32((uint32_t)strlen(seqs[q]));
				refs.u64(a_col);
				refs.u32(id++);
			}
		}
	}
	F.eof = (F.eof + 7) & ~7ull;
	const uint64_t a_seq_data = F.append(refs.b.data(), refs.size());

	// ---- /similarity_matrix --------------------------------------------------------------
	uint64_t a_mat = 0;
	const bool tri = triangular != 0;
	if (!chunked) {
		std::vector<int32_t> full(dim * dim);
		for (size_t r = 0; r < dim; r++)
			for (size_t c = 0; c < dim; c++)
				full[r * dim + c] = cell_at(matrix, tri, dim, r, c);
		F.eof = (F.eof + 7) & ~7ull;
		a_mat = F.append(full.data(), full.size() * 4);
	} else {
		const size_t nc = (dim + cdim - 1) / cdim;
		const size_t raw_bytes = cdim * cdim * 4;
		std::vector<ChunkRec> recs;
		recs.reserve(nc * nc);
		unsigned hw = std::thread::hardware_concurrency();
		const size_t total = nc * nc; // chunks in row-major order (the order they are written in)
		const size_t nthreads = std::max<size_t>(1, std::min<size_t>({ (size_t)(hw ? hw : 4), 16, total }));
		// Batches of `nthreads` chunks across chunk-row boundaries, two slot sets: batch b+1 is
		// gathered + deflated while the calling thread appends batch b to the file, in order.
		struct Slot {
			std::vector<int32_t> raw;
			std::vector<uint8_t> z;
			size_t zlen = 0;
			bool err = false;
		};
		std::vector<Slot> slots(deflate ? std::min(total, 2 * nthreads) : 0);
		for (auto &s : slots) {
			s.raw.resize(cdim * cdim);
			s.z.resize(compressBound(raw_bytes));
		}
		// gathers chunk `idx` into raw[cdim * cdim] (edge chunks are stored whole, padded with 0)
		auto gather = [&](const size_t idx, int32_t *raw) {
			const size_t cr = idx / nc, cc = idx % nc;
			const uint64_t R0 = cr * cdim, C0 = cc * cdim;
			if (tri && cc > cr) {
				// block strictly above the diagonal: element (R, C) is packed-row C, column R
				// (src/io/output.c:83), i.e. the transpose of contiguous runs -> 32 x 32 tiles
				const size_t h = (size_t)std::min<uint64_t>(cdim, dim - R0);
				const size_t w = (size_t)std::min<uint64_t>(cdim, dim - C0);
				if (h < cdim || w < cdim)
					memset(raw, 0, raw_bytes);
				for (size_t c0 = 0; c0 < w; c0 += 32)
					for (size_t r0 = 0; r0 < h; r0 += 32) {
						const size_t ce = std::min(w, c0 + 32), re = std::min(h, r0 + 32);
						for (size_t c = c0; c < ce; c++) {
							const uint64_t C = C0 + c;
							const int32_t *src = matrix + C * (C - 1) / 2 + R0;
							for (size_t r = r0; r < re; r++)
								raw[r * cdim + c] = src[r];
						}
					}
			} else
			for (size_t r = 0; r < cdim; r++) {
				int32_t *dst = raw + r * cdim;
				const uint64_t R = R0 + r;
				if (R >= dim) { // edge chunks are stored whole, padded with 0
					memset(dst, 0, cdim * 4);
					continue;
				}
				const size_t w = (size_t)std::min<uint64_t>(cdim, dim - C0);
				if (!tri) {
					memcpy(dst, matrix + R * dim + C0, w * 4);
				} else if (cc < cr) { // strictly below the diagonal: a run of packed row R
					memcpy(dst, matrix + R * (R - 1) / 2 + C0, w * 4);
				} else {
					for (size_t c = 0; c < w; c++)
						dst[c] = cell_at(matrix, true, dim, R, C0 + c);
				}
				if (w < cdim)
					memset(dst + w, 0, (cdim - w) * 4);
			}
		};
		auto work = [&](const size_t idx, Slot &s) {
			gather(idx, s.raw.data());
			if (deflate) {
				uLongf zl = (uLongf)s.z.size();
				s.err = compress2(s.z.data(), &zl, (const Bytef *)s.raw.data(), (uLong)raw_bytes,
						  (int)compression) != Z_OK;
				s.zlen = zl;
			}
		};
		if (!deflate) {
			// Uncompressed chunks have known file addresses: the workers gather a chunk into their own
			// buffer and pwrite it, independently and in any order (no batch barrier, no serial append,
			// no zero-filled slot sets: a 1 GB matrix into tmpfs on 8 cores went from 0.55 to 1.6 GB/s).
			fflush(F.f);
			const int fd = fileno(F.f);
			const uint64_t base = F.eof;
			// First choice: gather straight into a shared mapping of the chunk area (the scores are
			// written once, into the page cache); fallback: a buffer per worker + pwrite.
			const uint64_t page = (uint64_t)sysconf(_SC_PAGESIZE);
			const uint64_t map_off = base / page * page;
			const size_t map_len = (size_t)(base - map_off + total * raw_bytes);
			unsigned char *map = nullptr;
			if (ftruncate(fd, (off_t)(base + total * raw_bytes)) == 0) {
				void *m = mmap(nullptr, map_len, PROT_READ | PROT_WRITE, MAP_SHARED, fd, (off_t)map_off);
				if (m != MAP_FAILED)
					map = static_cast<unsigned char *>(m);
			}
			std::atomic<size_t> next{ 0 }, done{ 0 };
			std::atomic<bool> bad{ false };
			auto worker = [&]() {
				if (map) {
					for (;;) {
						const size_t idx = next.fetch_add(1);
						if (idx >= total)
							break;
						gather(idx, reinterpret_cast<int32_t *>(map + (base - map_off) + idx * raw_bytes));
						done.fetch_add(1);
					}
					return;
				}
				int32_t *raw = new (std::nothrow) int32_t[cdim * cdim];
				if (!raw) {
					bad = true;
					return;
				}
				for (;;) {
					const size_t idx = next.fetch_add(1);
					if (idx >= total || bad)
						break;
					gather(idx, raw);
					const char *p = (const char *)raw;
					size_t left = raw_bytes;
					uint64_t off = base + idx * raw_bytes;
					while (left) {
						const ssize_t w = pwrite(fd, p, left, (off_t)off);
						if (w <= 0) {
							bad = true;
							break;
						}
						p += w;
						left -= (size_t)w;
						off += (uint64_t)w;
					}
					done.fetch_add(1);
				}
				delete[] raw;
			};
			std::vector<std::thread> th;
			for (size_t t = 0; t < nthreads; t++)
				th.emplace_back(worker);
			size_t reported = 0;
			while (done.load() < total && !bad.load()) { // the calling thread reports progress
				std::this_thread::sleep_for(std::chrono::milliseconds(5));
				const size_t rows = done.load() / nc;
				if (progress && rows > reported && rows < nc)
					progress((int64_t)(reported = rows), (int64_t)nc, user);
			}
			for (auto &t : th)
				t.join();
			if (map && munmap(map, map_len) != 0)
				bad = true;
			if (bad.load()) {
				fclose(F.f);
				return SAB_ERR_NOMEM;
			}
			if (progress)
				progress((int64_t)nc, (int64_t)nc, user);
			for (size_t idx = 0; idx < total; idx++)
				recs.push_back({ base + idx * raw_bytes, (uint32_t)raw_bytes, (idx / nc) * cdim, (idx % nc) * cdim });
			F.eof = base + total * raw_bytes;
		}
		auto launch = [&](const size_t base, const size_t set, std::vector<std::thread> &th) {
			const size_t nb = std::min(nthreads, total - base);
			for (size_t t = 0; t < nb; t++)
				th.emplace_back([&work, &slots, base, set, t, nthreads]() {
					work(base + t, slots[(set * nthreads + t) % slots.size()]);
				});
		};
		std::vector<std::thread> cur, nxt;
		size_t set = 0;
		if (deflate)
			launch(0, set, cur);
		for (size_t base = 0; deflate && base < total;) {
			const size_t nb = std::min(nthreads, total - base);
			if (base + nb < total && F.ok)
				launch(base + nb, set ^ 1, nxt);
			for (auto &t : cur)
				t.join();
			cur.clear();
			for (size_t t = 0; t < nb && F.ok; t++) {
				Slot &s = slots[(set * nthreads + t) % slots.size()];
				if (s.err) {
					for (auto &x : nxt)
						x.join();
					fclose(F.f);
					return SAB_ERR_NOMEM;
				}
				const size_t idx = base + t;
				const void *p = deflate ? (const void *)s.z.data() : (const void *)s.raw.data();
				const size_t n = deflate ? s.zlen : raw_bytes;
				const uint64_t a = F.append(p, n);
				recs.push_back({ a, (uint32_t)n, (idx / nc) * cdim, (idx % nc) * cdim });
				if (progress && (idx + 1) % nc == 0)
					progress((int64_t)((idx + 1) / nc), (int64_t)nc, user);
			}
			cur.swap(nxt);
			base += nb;
			set ^= 1;
		}
		for (auto &t : cur) // only reached with threads left after a write error
			t.join();
		a_mat = write_chunk_tree(F, recs, cdim);
	}

	// ---- the metadata block ----------------------------------------------------------------
	{
		Buf sb;
		const uint8_t sig[8] = { 0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n' };
		sb.raw(sig, 8);
		sb.u8(0); // superblock version
		sb.u8(0); // free-space storage version
		sb.u8(0); // root group symbol table entry version
		sb.u8(0);
		sb.u8(0); // shared header message format version
		sb.u8(8); // size of offsets
		sb.u8(8); // size of lengths
		sb.u8(0);
		sb.u16(kGroupLeafK);
		sb.u16(kGroupInternalK);
		sb.u32(0); // file consistency flags
		sb.u64(0); // base address
		sb.u64(UNDEF); // free-space info
		sb.u64(F.eof); // end of file address
		sb.u64(UNDEF); // driver information block
		// root group symbol table entry
		sb.u64(0); // link name offset
		sb.u64(a_root);
		sb.u32(1); // cache type: group (scratch = B-tree + heap)
		sb.u32(0);
		sb.u64(a_btree);
		sb.u64(a_heap);
		F.at(a_super, sb.b.data(), sb.size());

		Buf st;
		st.u64(a_btree);
		st.u64(a_heap);
		Buf rm;
		message(rm, 0x0011, st);
		const Buf rh = object_header(rm, 1);
		F.at(a_root, rh.b.data(), rh.size());

		Buf hp;
		hp.raw("HEAP", 4);
		hp.u8(0);
		hp.zero(3);
		hp.u64(heap_data.size());
		hp.u64(1); // H5HL_FREE_NULL: no free block
		hp.u64(a_heap_data);
		F.at(a_heap, hp.b.data(), hp.size());
		F.at(a_heap_data, heap_data.b.data(), heap_data.size());

		Buf bt;
		bt.raw("TREE", 4);
		bt.u8(0); // group node
		bt.u8(0); // leaf level
		bt.u16(1);
		bt.u64(UNDEF);
		bt.u64(UNDEF);
		bt.u64(0); // key 0: ""
		bt.u64(a_snod);
		bt.u64(n_mat); // key 1: largest name in the child
		bt.zero(group_node_size - bt.size());
		F.at(a_btree, bt.b.data(), bt.size());

		Buf sn;
		sn.raw("SNOD", 4);
		sn.u8(1);
		sn.u8(0);
		sn.u16(2);
		const uint64_t names[2] = { n_seq, n_mat }, objs[2] = { a_dset_seq, a_dset_mat };
		for (int k = 0; k < 2; k++) { // sorted by name: "sequences" < "similarity_matrix"
			sn.u64(names[k]);
			sn.u64(objs[k]);
			sn.u32(0);
			sn.u32(0);
			sn.zero(16);
		}
		sn.zero(snod_size - sn.size());
		F.at(a_snod, sn.b.data(), sn.size());

		const Buf sh = seq_header(a_seq_data);
		F.at(a_dset_seq, sh.b.data(), sh.size());
		const Buf mh = mat_header(a_mat);
		F.at(a_dset_mat, mh.b.data(), mh.size());
	}
	const bool ok = F.ok && fflush(F.f) == 0;
	fclose(F.f);
	return ok ? SAB_OK : SAB_ERR_ARG;
}
