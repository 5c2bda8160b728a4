// mesh_core.cuh -- per-lane data-path logic of the B200 chunk mesher (v3: analyse + expand).
//
// Everything in this header is written per *logical lane* against an explicit per-warp
// shared-memory struct; data crosses lanes only through that struct (or through values the
// caller hands in).  The CUDA kernels (mesh_kernel.cuh) call these functions with
// lane = threadIdx.x & 31, real shuffles and __syncwarp() between the phases; tests/emul/
// compiles the very same functions with g++ and runs the 32 lanes in a loop, so the
// bit-level logic can be checked on a machine without a GPU.  (That emulation is test
// infrastructure; the product library never runs it.)
//
// What is computed: mimas-meshgen/lib.rs:231-446 `mesh_for_chunk`
//   * six run-merging face passes (lib.rs:316-410) -> `intransparent`
//   * crossed-plant pass (lib.rs:412-440)          -> `transparent`
// re-formulated as two kernels.
// ANALYSE (one 128-thread CTA per chunk; "lane" below = the thread's job index in a phase):
//   A. lane <-> column (x, y): 16 block ids along z (one 16-byte load) -> 16-bit "blocky"
//      mask over z, plus SWAR id-equality masks against the (x, y-1) and (x-1, y) columns
//   B. visibility and run-start masks of all six passes as bit operations on z-masks
//   C. popc of the start masks -> quads per (pass, c1) group, warp scan -> group offsets
//   D. 16x16 bit-matrix transposes (in registers) turn the z-masks into masks over the
//      merge axis, one 16-bit word per row (pass, c1, c2) in the reference's emission order;
//      every run start becomes one 32-bit quad descriptor (block id, start, row, pass, length)
//      in the chunk's descriptor slot; crossed plants become plant descriptors
// EXPAND (one warp per tile of 32 quads of the packed stream):
//   E. exclusive scan of the per-chunk counts (shared memory, redundantly per CTA), quad ->
//      chunk by binary search, descriptor -> 216 bytes into a staging tile that a bulk async
//      copy (TMA) stores to the packed vertex stream.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define MM_HD __host__ __device__ __forceinline__
#else
#define MM_HD static inline
struct alignas(8) uint2 {
	unsigned int x, y;
};
#endif

namespace mm {

typedef unsigned long long u64;

struct Range {  // = mm_range (include/mimas_meshgen_cuda.h)
	u64 first_vertex;
	uint32_t vertex_count;
	uint32_t reserved;
};

constexpr int kLanes = 32;
constexpr int kPasses = 6;             // lib.rs:316-410
constexpr int kGroups = 96;            // (pass, c1) groups of 16 rows, in emission order
constexpr int kTileQuads = 32;         // quads expanded per staging tile (one per lane)
constexpr int kQuadBytes = 216;        // 6 vertices x 36 B
constexpr int kQuadWords = 54;
constexpr int kPlantQuads = 4;         // 2 x rpush_face_bidi! = 24 vertices = 4 quad slots
constexpr int kStageBytes = kTileQuads * kQuadBytes;  // 6912, a multiple of 16
constexpr int kBzStride = 18;          // Bzp is a 18 x 18 array: the 16 x 16 columns plus a zero border
// Descriptor slots of one chunk (32-bit words): the most quads the six passes can emit
// (3-D checkerboard of two textures: 6 x 256 rows x 8... bounded by 6 x 4096 / 2) and the most plants.
constexpr uint32_t kDescStrideI = 12288;
constexpr uint32_t kDescStrideT = 4096;

// Block-table entry as the kernels see it (built from mm_block_entry on the host):
//   x = id_sides | id_top << 16   (for a crossed block: x = crossed texture id)
//   y = id_bottom | kind << 16    (kind: 0 none, 1 blocky, 2 crossed)
struct TabEntry {
	uint32_t x, y;
};
// Per-id flag byte derived from the table (build_flags): bit 0 = blocky, bit 1 = crossed, bits 2..4 =
// this blocky id shares its sides / top / bottom texture with a *different* blocky id.  Where no cell
// of a chunk has such a bit, "different block ids" already means "different textures" and the run
// merging never needs to look a texture up.
constexpr uint32_t kFlagBlocky = 1u, kFlagCrossed = 2u, kFlagAmbSides = 4u, kFlagAmbTop = 8u, kFlagAmbBottom = 16u;
constexpr uint32_t kFlagAmbAny = kFlagAmbSides | kFlagAmbTop | kFlagAmbBottom;

// 16 block ids of one column (x, y), z = 0..15: byte z&3 of word z>>2 (map.rs:83-86).
struct Col {
	uint32_t w[4];
};

// Shared memory of one analysing CTA (= one chunk in flight).  Plain struct so that host
// emulation can allocate it.
struct alignas(128) ChunkSmem {
	// the chunk itself, filled by one bulk async copy (TMA)
	alignas(128) unsigned char raw[4096];
	// phase D: wl, the list of non-empty 32-bit words of S (two rows each) with their quad offsets;
	// plants: Cz, the crossed mask over z per column
	alignas(16) uint32_t u[768];
	// run-start masks over z per column: passes 1,3,4,6 (p = 0,2,3,5) at [p][x*16 + y],
	// passes 2,5 (p = 1,4) at [p][y*16 + x]; every 16-entry block [p][c1*16 ..] is one
	// (pass, c1) group whose 16 x 16 bit transpose gives the masks over the merge axis
	alignas(16) uint16_t S[kPasses * 256];
	alignas(16) uint16_t By[256];              // blocky mask over y for row (x, z): [x*16 + z]
	alignas(16) uint16_t Bx[256];              // blocky mask over x for row (y, z): [y*16 + z]
	alignas(16) uint16_t Bzp[kBzStride * kBzStride + 4];  // blocky mask over z per column, zero border
	alignas(16) uint32_t gcnt[kGroups];        // per group: quads | non-empty words << 16; then the exclusive prefix of that
	// halo mode only (an extension, see halo_row): blocky masks of the touching layers of the z-, z+ neighbour
	// chunks over y per x (Hz[x], Hz[16 + x]) and of the x-, x+, y-, y+ neighbours over the merge axis per z
	// (Hm[0..15] x- over y, Hm[16..31] x+ over y, Hm[32..47] y- over x, Hm[48..63] y+ over x)
	alignas(16) uint16_t Hz[32];
	alignas(16) uint16_t Hm[64];
	uint32_t amb[2];                           // OR of the flag bytes of the chunk's cells (double buffered)
	uint32_t ticket[2];                        // chunk index of this / the next round (double buffered)
	alignas(8) unsigned long long mbar;        // mbarrier of the chunk load (device only)
};

MM_HD unsigned char *smem_raw(ChunkSmem &s) { return s.raw; }
MM_HD uint16_t *smem_Bzp(ChunkSmem &s) { return s.Bzp; }
MM_HD const uint16_t *smem_Bzp(const ChunkSmem &s) { return s.Bzp; }
MM_HD uint16_t *smem_Cz(ChunkSmem &s) { return reinterpret_cast<uint16_t *>(s.u); }   // crossed mask over z per column
MM_HD const uint16_t *smem_Cz(const ChunkSmem &s) { return reinterpret_cast<const uint16_t *>(s.u); }
// wl entry: index of a non-empty word of S (m*8 + k, 10 bits) | quads of the chunk before it << 10
MM_HD uint32_t *smem_wl(ChunkSmem &s) { return s.u; }
MM_HD const uint32_t *smem_wl(const ChunkSmem &s) { return s.u; }

MM_HD int popc32(uint32_t v) {
#if defined(__CUDA_ARCH__)
	return __popc(v);
#else
	return __builtin_popcount(v);
#endif
}
MM_HD int ffs32(uint32_t v) {  // 1-based, 0 if v == 0
#if defined(__CUDA_ARCH__)
	return __ffs((int)v);
#else
	return __builtin_ffs((int)v);
#endif
}
MM_HD uint32_t funnel_r1(uint32_t lo, uint32_t hi) {  // (hi:lo) >> 1, low word
#if defined(__CUDA_ARCH__)
	return __funnelshift_r(lo, hi, 1);
#else
	return (lo >> 1) | (hi << 31);
#endif
}
MM_HD uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t sel) {
#if defined(__CUDA_ARCH__)
	return __byte_perm(a, b, sel);
#else
	const u64 v = ((u64)b << 32) | a;
	uint32_t r = 0;
	for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
	return r;
#endif
}
// IEEE binary32 add / sub, round to nearest even, never contracted into an FMA --
// the Rust reference has no fused operations.
MM_HD float fadd(float a, float b) {
#if defined(__CUDA_ARCH__)
	return __fadd_rn(a, b);
#else
	volatile float r = a + b;
	return r;
#endif
}
MM_HD float fsub(float a, float b) {
#if defined(__CUDA_ARCH__)
	return __fsub_rn(a, b);
#else
	volatile float r = a - b;
	return r;
#endif
}
MM_HD float i64_to_f32(long long v) {  // Rust `isize as f32`: round to nearest even
#if defined(__CUDA_ARCH__)
	return __ll2float_rn(v);
#else
	return (float)v;
#endif
}
MM_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
	return __float_as_uint(f);
#else
	union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
// mimas-meshgen/lib.rs:123-131 `zsig` (NOT signum: +-0 -> +0)
MM_HD float zsig(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }

// Flag bytes of a 256-entry table (host side of mm_set_block_table; also used by the emulation).
static inline void build_flags(const TabEntry *tab, uint8_t *flags) {
	for (int i = 0; i < 256; i++) {
		const uint32_t kind = tab[i].y >> 16;
		uint32_t f = kind == 1 ? kFlagBlocky : kind == 2 ? kFlagCrossed : 0u;
		if (kind == 1)
			for (int j = 0; j < 256; j++) {
				if (j == i || (tab[j].y >> 16) != 1) continue;
				if ((tab[i].x & 0xffffu) == (tab[j].x & 0xffffu)) f |= kFlagAmbSides;
				if ((tab[i].x >> 16) == (tab[j].x >> 16)) f |= kFlagAmbTop;
				if ((tab[i].y & 0xffffu) == (tab[j].y & 0xffffu)) f |= kFlagAmbBottom;
			}
		flags[i] = (uint8_t)f;
	}
}

// ---- warp set-up ---------------------------------------------------------------------
// Once per CTA (lane = 0..31 of one warp): the zero border of Bzp (a neighbour outside the chunk never blocks,
// lib.rs:277-280); the interior is rewritten by phase A for every chunk.
MM_HD void init_warp_smem(ChunkSmem &s, int lane) {
	uint16_t *B = smem_Bzp(s);
	if (lane < kBzStride) {
		B[lane] = 0;
		B[(kBzStride - 1) * kBzStride + lane] = 0;
	}
	if (lane < 16) {
		B[(lane + 1) * kBzStride] = 0;
		B[(lane + 1) * kBzStride + kBzStride - 1] = 0;
	}
}

// 16 block ids of one column from the chunk in shared memory (16-byte aligned)
MM_HD Col load_col(const uint8_t *p) {
	Col c;
#if defined(__CUDA_ARCH__)
	const uint4 v = *reinterpret_cast<const uint4 *>(p);
	c.w[0] = v.x; c.w[1] = v.y; c.w[2] = v.z; c.w[3] = v.w;
#else
	memcpy(c.w, p, 16);
#endif
	return c;
}

// ---- phase A: one column (x, y) ------------------------------------------------------
// Blocky mask over z (bit z <=> cell (x, y, z) is drawn as a cube, i.e. get_bl_tex_ids is
// Some, lib.rs:105-109) in the return value; `orf` receives the OR of the column's flag bytes.
// flags = 256-entry byte table (see kFlagBlocky).
MM_HD uint32_t col_flags(const uint8_t *flags, const Col &c, uint32_t &orf) {
	uint32_t acc = 0, o = 0;
#pragma unroll
	for (int k = 0; k < 4; k++) {
#pragma unroll
		for (int b = 0; b < 4; b += 2) {
			const uint32_t f0 = flags[(c.w[k] >> (8 * b)) & 0xffu], f1 = flags[(c.w[k] >> (8 * b + 8)) & 0xffu];
			acc = funnel_r1(acc, f0);  // bit 0 of f enters at bit 31
			acc = funnel_r1(acc, f1);
			o |= f0 | f1;
		}
	}
	orf = o;
	return acc >> 16;
}

// Bit z set <=> the two columns hold the same block id at z (equal ids => equal textures
// for every face; different ids are settled exactly by fix_equal()).
MM_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
	return __umulhi(a, b);
#else
	return (uint32_t)(((u64)a * (u64)b) >> 32);
#endif
}
MM_HD uint32_t funnel_r4(uint32_t lo, uint32_t hi) {  // (hi:lo) >> 4, low word
#if defined(__CUDA_ARCH__)
	return __funnelshift_r(lo, hi, 4);
#else
	return (lo >> 4) | (hi << 28);
#endif
}
MM_HD uint32_t eq_mask16(const Col &a, const Col &b) {
	uint32_t acc = 0;
#pragma unroll
	for (int k = 0; k < 4; k++) {
		const uint32_t v = a.w[k] ^ b.w[k];
		const uint32_t t = ((v & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v;  // bit 7 of a byte: ids differ
		const uint32_t u = ~t & 0x80808080u;                       // bit 7 of a byte: ids equal
		// bits 7,15,23,31 -> bits 32..35 of the 64-bit product (multiplier 2^25 + 2^18 + 2^11 + 2^4; all
		// partial products land on distinct bits, so nothing carries); the nibble enters acc from the top
		acc = funnel_r4(acc, mulhi32(u, 0x02040810u));
	}
	return acc >> 16;
}

// Crossed mask over z of one column (slow path, only for columns whose flag sum says so).
MM_HD uint32_t col_crossed(const uint8_t *flags, const Col &c) {
	uint32_t m = 0;
#pragma unroll
	for (int z = 0; z < 16; z++)
		m |= ((flags[(c.w[z >> 2] >> (8 * (z & 3))) & 0xffu] >> 1) & 1u) << z;
	return m;
}

// ---- phase B: visibility and run starts of one column, all six passes -----------------
// lib.rs:273-293: a face is visible iff the cell is blocky and the neighbour in the pass
// direction is outside the chunk or not blocky.  A run (lib.rs:244-271) starts at a cell
// iff its face is visible and the previous cell along the merge axis is not visible or
// has another texture for that face.
//
// Exact texture comparison for the (rare) pairs of visible faces with different block ids.
// `d` = z bits to settle, cells (x, y, z) vs (px, py, z); face: 0 sides, 1 top, 2 bottom.
MM_HD uint32_t fix_equal(const TabEntry *tab, const uint8_t *chunk, uint32_t d, int x, int y, int px, int py,
                         int face) {
	uint32_t eq = 0;
	while (d) {
		const int z = ffs32(d) - 1;
		d &= d - 1u;
		const TabEntry a = tab[chunk[x * 256 + y * 16 + z]], b = tab[chunk[px * 256 + py * 16 + z]];
		const uint32_t ta = face == 0 ? (a.x & 0xffffu) : face == 1 ? (a.x >> 16) : (a.y & 0xffffu);
		const uint32_t tb = face == 0 ? (b.x & 0xffffu) : face == 1 ? (b.x >> 16) : (b.y & 0xffffu);
		eq |= (ta == tb ? 1u : 0u) << z;
	}
	return eq;
}

// `amb` = OR of the flag bytes of every cell of the chunk.
// HALO: the border of Bzp and Hz hold the neighbour chunks' touching layers (halo_row); they occlude
// faces like cells of the chunk do, but a run never continues from them (a row starts at 0, lib.rs:294-313).
template <bool HALO>
MM_HD void col_starts(ChunkSmem &s, const TabEntry *tab, const uint8_t *chunk, int x, int y, uint32_t amb) {
	const uint16_t *B = smem_Bzp(s) + (x + 1) * kBzStride + (y + 1);
	const uint32_t b = B[0];
	const uint32_t bym = B[-1], byp = B[1], bxm = B[-kBzStride], bxp = B[kBzStride];
	const uint32_t bmm = B[-kBzStride - 1], bpm = B[kBzStride - 1], bmp = B[-kBzStride + 1];
	const int col = x * 16 + y;
	// previous cell along the merge axis; cells below z = 0 / above z = 15 of this and of the previous column
	uint32_t py = bym, px = bxm, zlo = 0, zhi = 0, pzlo = 0, pzhi = 0;
	if (HALO) {
		const uint32_t hm = s.Hz[x], hp = s.Hz[16 + x];
		py = y > 0 ? bym : 0u;
		px = x > 0 ? bxm : 0u;
		zlo = (hm >> y) & 1u;
		zhi = ((hp >> y) & 1u) << 15;
		pzlo = ((hm << 1) >> y) & 1u;
		pzhi = (((hp << 1) >> y) & 1u) << 15;
	}
	// own faces
	const uint32_t V1 = b & ~((b << 1) | zlo), V4 = b & ~((b >> 1) | zhi);   // [0,0,-1] lib.rs:317, [0,0,1] lib.rs:365
	const uint32_t V3 = b & ~bxm, V6 = b & ~bxp;             // [-1,0,0] lib.rs:349, [1,0,0] lib.rs:397
	const uint32_t V2 = b & ~bym, V5 = b & ~byp;             // [0,-1,0] lib.rs:333, [0,1,0] lib.rs:381
	// the same faces of the previous cell along the merge axis: (x, y-1) for the y-inner
	// passes 1,3,4,6 and (x-1, y) for the x-inner passes 2,5 (zero border: no previous cell)
	const uint32_t Q1 = V1 & py & ~((py << 1) | pzlo), Q4 = V4 & py & ~((py >> 1) | pzhi);
	const uint32_t Q3 = V3 & py & ~bmm, Q6 = V6 & py & ~bpm;
	const uint32_t Q2 = V2 & px & ~bmm, Q5 = V5 & px & ~bmp;
	// Runs merge on equal textures; equal block ids are equal textures, and the id-equality masks
	// against the previous column are only worked out where two visible faces meet at all.
	const uint32_t qy = Q1 | Q3 | Q4 | Q6, qx = Q2 | Q5;
	uint32_t ey = 0, ex = 0;
	if (qy | qx) {
		const Col own = load_col(chunk + col * 16);
		if (qy) ey = eq_mask16(own, load_col(chunk + (col - 1) * 16));    // (x, y-1): exists, its blocky mask is not empty
		if (qx) ex = eq_mask16(own, load_col(chunk + (col - 16) * 16));   // (x-1, y)
	}
	uint32_t e1 = ey, e3 = ey, e4 = ey, e6 = ey, e2 = ex, e5 = ex;
	// rare: visible neighbours with different ids in a chunk where different ids may share a texture
	if ((amb & kFlagAmbAny) && (((Q1 | Q3 | Q4 | Q6) & ~ey) | ((Q2 | Q5) & ~ex))) {
		if (amb & kFlagAmbBottom) e1 |= fix_equal(tab, chunk, Q1 & ~ey, x, y, x, y - 1, 2);  // id_bottom, lib.rs:329
		if (amb & kFlagAmbTop) e4 |= fix_equal(tab, chunk, Q4 & ~ey, x, y, x, y - 1, 1);     // id_top,    lib.rs:377
		if (amb & kFlagAmbSides) {
			e3 |= fix_equal(tab, chunk, Q3 & ~ey, x, y, x, y - 1, 0);  // id_sides,  lib.rs:361
			e6 |= fix_equal(tab, chunk, Q6 & ~ey, x, y, x, y - 1, 0);  // id_sides,  lib.rs:409
			e2 |= fix_equal(tab, chunk, Q2 & ~ex, x, y, x - 1, y, 0);  // id_sides,  lib.rs:345
			e5 |= fix_equal(tab, chunk, Q5 & ~ex, x, y, x - 1, y, 0);  // id_sides,  lib.rs:393
		}
	}
	const int row = y * 16 + x;
	s.S[0 * 256 + col] = (uint16_t)(V1 & ~(Q1 & e1));
	s.S[1 * 256 + row] = (uint16_t)(V2 & ~(Q2 & e2));
	s.S[2 * 256 + col] = (uint16_t)(V3 & ~(Q3 & e3));
	s.S[3 * 256 + col] = (uint16_t)(V4 & ~(Q4 & e4));
	s.S[4 * 256 + row] = (uint16_t)(V5 & ~(Q5 & e5));
	s.S[5 * 256 + col] = (uint16_t)(V6 & ~(Q6 & e6));
}

// ---- phase C: 16 x 16 bit-matrix transpose in registers --------------------------------
// w[k] holds rows 2k (low half) and 2k+1 (high half); bit c of row r <-> bit r of row c.
MM_HD void transpose16(uint32_t w[8]) {
	// 8 x 8 blocks: byte 1 of rows 0..7 <-> byte 0 of rows 8..15
#pragma unroll
	for (int k = 0; k < 4; k++) {
		const uint32_t a = w[k], b = w[k + 4];
		w[k] = byte_perm(a, b, 0x6240);      // a.b0, b.b0, a.b2, b.b2
		w[k + 4] = byte_perm(a, b, 0x7351);  // a.b1, b.b1, a.b3, b.b3
	}
	// 4 x 4 blocks: rows r and r+4
#pragma unroll
	for (int k = 0; k < 8; k++) {
		if (k & 2) continue;
		const uint32_t t = ((w[k] >> 4) ^ w[k + 2]) & 0x0f0f0f0fu;
		w[k + 2] ^= t;
		w[k] ^= t << 4;
	}
	// 2 x 2 blocks: rows r and r+2
#pragma unroll
	for (int k = 0; k < 8; k += 2) {
		const uint32_t t = ((w[k] >> 2) ^ w[k + 1]) & 0x33333333u;
		w[k + 1] ^= t;
		w[k] ^= t << 2;
	}
	// 1 x 1: rows 2k and 2k+1 live in one word: bit 2j+1 <-> bit 16+2j
#pragma unroll
	for (int k = 0; k < 8; k++) {
		const uint32_t t = ((w[k] >> 15) ^ w[k]) & 0x0000aaaau;
		w[k] ^= t ^ (t << 15);
	}
}


// Blocky masks over the merge axes, from the z-masks: job 0..15 -> By rows of x = job
// (bits over y), job 16..31 -> Bx rows of y = job - 16 (bits over x).
MM_HD void transpose_blocky(ChunkSmem &s, int job) {
	uint32_t w[8];
	const bool over_y = job < 16;
	const int f = job & 15;
#pragma unroll
	for (int k = 0; k < 8; k++) {
		const int r0 = 2 * k, r1 = 2 * k + 1;
		const uint16_t *B = smem_Bzp(s);
		const uint32_t lo = over_y ? B[(f + 1) * kBzStride + r0 + 1] : B[(r0 + 1) * kBzStride + f + 1];
		const uint32_t hi = over_y ? B[(f + 1) * kBzStride + r1 + 1] : B[(r1 + 1) * kBzStride + f + 1];
		w[k] = lo | (hi << 16);
	}
	transpose16(w);
	uint32_t *p = reinterpret_cast<uint32_t *>((over_y ? s.By : s.Bx) + f * 16);
#pragma unroll
	for (int k = 0; k < 8; k++) p[k] = w[k];
}

// ---- halo mode (extension; the reference has no halo: lib.rs:277-280) -----------------------------
// Per chunk 6 slabs of 256 block ids, the layers of the face-neighbour chunks that touch this chunk, in the
// order x-, x+, y-, y+, z-, z+; x slabs indexed y*16 + z, y slabs x*16 + z, z slabs x*16 + y (the same
// layout as oracle/meshgen_oracle.c blocked()).  Job t = 0..95 takes the 16 ids of row t & 15 of slab
// t >> 4 (`c`) to a blocky mask: the x / y slabs give masks over z for the border columns of Bzp, the
// z slabs masks over y per x.
MM_HD void halo_row(ChunkSmem &s, const uint8_t *flags, const Col &c, int t) {
	uint32_t orf;
	const uint16_t m = (uint16_t)col_flags(flags, c, orf);
	const int slab = t >> 4, r = t & 15;
	uint16_t *B = smem_Bzp(s);
	if (slab == 0) B[r + 1] = m;
	else if (slab == 1) B[(kBzStride - 1) * kBzStride + r + 1] = m;
	else if (slab == 2) B[(r + 1) * kBzStride] = m;
	else if (slab == 3) B[(r + 1) * kBzStride + kBzStride - 1] = m;
	else s.Hz[(slab - 4) * 16 + r] = m;
}
// Job 0..3: the border masks of the x-, x+, y-, y+ layer (over z per y resp. x) transposed to masks over
// the merge axis per z, as phase D needs them.
MM_HD void transpose_halo(ChunkSmem &s, int job) {
	uint32_t w[8];
	const uint16_t *B = smem_Bzp(s);
#pragma unroll
	for (int k = 0; k < 8; k++) {
		uint32_t lo, hi;
		if (job < 2) {
			const int x = job == 0 ? 0 : kBzStride - 1;
			lo = B[x * kBzStride + 2 * k + 1];
			hi = B[x * kBzStride + 2 * k + 2];
		} else {
			const int y = job == 2 ? 0 : kBzStride - 1;
			lo = B[(2 * k + 1) * kBzStride + y];
			hi = B[(2 * k + 2) * kBzStride + y];
		}
		w[k] = lo | (hi << 16);
	}
	transpose16(w);
	uint32_t *p = reinterpret_cast<uint32_t *>(s.Hm + job * 16);
#pragma unroll
	for (int k = 0; k < 8; k++) p[k] = w[k];
}

// Transposes the 16-entry block at S[m*16 ..] in place: from masks over z per column to masks
// over the merge axis per row c2, i.e. word k of the block = rows 2k (low half) and 2k+1 (high
// half) of group m = pass*16 + c1 in emission order.  Returns quads | non-empty words << 16.
MM_HD uint32_t transpose_group(ChunkSmem &s, int m) {
	uint32_t *p = reinterpret_cast<uint32_t *>(s.S + m * 16);
	uint32_t w[8];
	uint32_t any = 0;
#pragma unroll
	for (int k = 0; k < 8; k++) {
		w[k] = p[k];
		any |= w[k];
	}
	if (any == 0) return 0;
	transpose16(w);
	uint32_t r = 0;
#pragma unroll
	for (int k = 0; k < 8; k++) {
		p[k] = w[k];
		r += (uint32_t)popc32(w[k]) + (w[k] ? 0x10000u : 0u);
	}
	return r;
}

// Appends the non-empty words of group m to wl.  `pre` = quads | words << 16 of the groups before m.
struct alignas(16) Word4 {
	uint32_t v[4];
};
MM_HD void list_group_words(ChunkSmem &s, int m, uint32_t pre) {
	// the group's eight words as two 16-byte loads: word by word the lanes (32 bytes apart) collide eight-fold
	const Word4 lo = *reinterpret_cast<const Word4 *>(s.S + m * 16), hi = *reinterpret_cast<const Word4 *>(s.S + m * 16 + 8);
	uint32_t *wl = smem_wl(s);
	uint32_t o = pre >> 16, q = pre & 0xffffu;
#pragma unroll
	for (int k = 0; k < 8; k++) {
		const uint32_t w = k < 4 ? lo.v[k] : hi.v[k - 4];
		if (w) {
			wl[o++] = (uint32_t)(m * 8 + k) | (q << 10);
			q += (uint32_t)popc32(w);
		}
	}
}

// ---- phase D: quad descriptors --------------------------------------------------------
// One 32-bit word per quad of the six face passes:
//   bits 0-7 block id of the run's first cell | 8-11 start (inner coordinate) | 12-15 c2 |
//   16-19 c1 | 20-22 pass (0..5) | 23-27 run length (1..16)
// Writes the descriptors of the runs starting in word `widx` of S (rows c2 = 2k and 2k+1 of
// group m, widx = m*8 + k; `word` = that word, non-zero) in emission order (rows ascending,
// runs ascending, lib.rs:294-313) to dst[0..popc(word)).  A run (lib.rs:244-271) ends where
// the next one starts, where the face stops being visible, or at the end of the row.
template <bool HALO>
MM_HD void emit_word_descs(const ChunkSmem &s, uint32_t widx, uint32_t word, uint32_t *dst) {
	const int m = (int)(widx >> 3), c2lo = (int)(widx & 7u) * 2;
	const int p = m >> 4, c1 = m & 15;
	const bool xin = (p == 1 || p == 4);
	// visibility over the merge axis of row (c1, c2) = blocky & ~blocky of the neighbour row in the
	// pass direction (SURVEY.md 8a pass table), for the two rows of the word at once
	const uint16_t *Bm = (xin ? s.Bx : s.By) + c1 * 16 + c2lo;
	const uint32_t b = *reinterpret_cast<const uint32_t *>(Bm);
	uint32_t nb;
	// beyond the chunk: nothing (lib.rs:277-280), or the neighbour chunk's layer in halo mode
	uint32_t out = 0;
	if (HALO) {
		if (p == 0) out = s.Hz[c1];
		else if (p == 3) out = (uint32_t)s.Hz[16 + c1] << 16;
		else out = *reinterpret_cast<const uint32_t *>(s.Hm + (p == 2 ? 0 : p == 5 ? 16 : p == 1 ? 32 : 48) + c2lo);
	}
	if (p == 0) nb = (b << 16) | (c2lo > 0 ? (uint32_t)Bm[-1] : out);                  // [0,0,-1]: row c2-1
	else if (p == 3) nb = (b >> 16) | (c2lo < 14 ? (uint32_t)Bm[2] << 16 : out);       // [0,0,1]:  row c2+1
	else if (p == 1 || p == 2) nb = c1 > 0 ? *reinterpret_cast<const uint32_t *>(Bm - 16) : out;  // [0,-1,0] / [-1,0,0]
	else nb = c1 < 15 ? *reinterpret_cast<const uint32_t *>(Bm + 16) : out;            // [0,1,0] / [1,0,0]
	const uint32_t T = (word | ~(b & ~nb) | 0x10000u) >> 1;  // run boundaries: start | not visible | end of the low row
	const uint32_t base = ((uint32_t)m << 16) | ((uint32_t)c2lo << 12);
	const int cell0 = (xin ? c1 * 16 : c1 * 256) + c2lo;   // + st*256 (x-inner) or st*16 (y-inner) + (c2 & 1)
	const int sh = xin ? 8 : 4;
	uint32_t rest = word;
	do {
		const int bit = ffs32(rest) - 1;  // = (c2 & 1) * 16 + st
		rest &= rest - 1u;
		const int len = ffs32((T >> bit) | (0x80000000u >> bit));  // .. | end of the high row
		const uint32_t id = s.raw[cell0 + ((bit & 15) << sh) + (bit >> 4)];
		*dst++ = base | ((uint32_t)bit << 8) | ((uint32_t)len << 23) | id;
	} while (rest);
}

// Crossed plants (lib.rs:412-440), x outer, y, z inner = linear cell order.  Cz[col] = crossed
// mask over z of column col = x*16 + y; this lane covers columns lane*8 .. lane*8+7.
MM_HD uint32_t count_plants8(const ChunkSmem &s, int lane) {
	const uint32_t *p = reinterpret_cast<const uint32_t *>(smem_Cz(s) + lane * 8);
	uint32_t c = 0;
#pragma unroll
	for (int k = 0; k < 4; k++) c += (uint32_t)popc32(p[k]);
	return c;
}
// Plant descriptor: bits 0-7 block id | 8-19 cell index.
MM_HD void emit_plant_descs(const ChunkSmem &s, int lane, uint32_t *dst) {
#pragma unroll 1
	for (int k = 0; k < 8; k++) {
		const int col = lane * 8 + k;
		uint32_t mask = smem_Cz(s)[col];
		while (mask) {
			const int z = ffs32(mask) - 1;
			mask &= mask - 1u;
			const int cell = col * 16 + z;
			*dst++ = (uint32_t)s.raw[cell] | ((uint32_t)cell << 8);
		}
	}
}

// ---- phase E: quad expansion -------------------------------------------------------------
// The argument list of rpush_face! / rpush_face_rev! / rpush_face_bidi!
// (lib.rs:139-181): base (x,y,z), position deltas (xsd,ysd,yd,zd), texture deltas
// summed as the macros sum them (us = xstd + ystd, vt = ytd + ztd), direction, tex.
struct QuadArgs {
	float x, y, z, xsd, ysd, yd, zd, us, vt;
	uint32_t tex;
	bool rev;
};

// Writes the 54 words of one quad.  `dst` must be 8-byte aligned.
MM_HD void write_quad(unsigned char *dst, const QuadArgs &a) {
	// corners, lib.rs:141-147
	const float p1x = fadd(a.x, a.xsd), p1y = fadd(a.y, a.ysd);             // P1 = (x+xsd, y+ysd, z)
	const float p2y = fadd(a.y, a.yd), p2z = fadd(a.z, a.zd);               // P2 = (x, y+yd, z+zd)
	const float p4y = fadd(fadd(a.y, a.yd), a.ysd);                         // P4 = (x+xsd, y+yd+ysd, z+zd)
	float nx, ny, nz;
	if (!a.rev) {  // sign![xsd, ysd + yd, zd]
		nx = zsig(a.xsd); ny = zsig(fadd(a.ysd, a.yd)); nz = zsig(a.zd);
	} else {       // sign![-xsd, -ysd - yd, -zd]
		nx = zsig(-a.xsd); ny = zsig(fsub(-a.ysd, a.yd)); nz = zsig(-a.zd);
	}
	// corner table: u, v, px, py, pz
	const float cu[4] = {0.0f, a.us, 0.0f, a.us};
	const float cv[4] = {0.0f, 0.0f, a.vt, a.vt};
	const float cx[4] = {a.x, p1x, a.x, p1x};
	const float cy[4] = {a.y, p1y, p2y, p4y};
	const float cz[4] = {a.z, a.z, p2z, p2z};
	// emission order: fwd P0,P1,P2,P1,P4,P2 (lib.rs:141-147); rev P2,P1,P0,P2,P4,P1 (lib.rs:152-158)
	uint32_t w[kQuadWords];
#pragma unroll
	for (int v = 0; v < 6; v++) {
		const int kf = (v == 0) ? 0 : (v == 1) ? 1 : (v == 2) ? 2 : (v == 3) ? 1 : (v == 4) ? 3 : 2;
		const int kr = (v == 0) ? 2 : (v == 1) ? 1 : (v == 2) ? 0 : (v == 3) ? 2 : (v == 4) ? 3 : 1;
		const bool rv = a.rev;
		w[v * 9 + 0] = a.tex;  // tex_ind:u16 | pad0 = 0
		w[v * 9 + 1] = f2u(rv ? cu[kr] : cu[kf]);
		w[v * 9 + 2] = f2u(rv ? cv[kr] : cv[kf]);
		w[v * 9 + 3] = f2u(rv ? cx[kr] : cx[kf]);
		w[v * 9 + 4] = f2u(rv ? cy[kr] : cy[kf]);
		w[v * 9 + 5] = f2u(rv ? cz[kr] : cz[kf]);
		w[v * 9 + 6] = f2u(nx);
		w[v * 9 + 7] = f2u(ny);
		w[v * 9 + 8] = f2u(nz);
	}
	uint2 *d2 = reinterpret_cast<uint2 *>(dst);
#pragma unroll
	for (int i = 0; i < kQuadWords / 2; i++) {
		uint2 t;
		t.x = w[2 * i];
		t.y = w[2 * i + 1];
		d2[i] = t;
	}
}

// One quad of the six face passes (pass table: SURVEY.md 8a A5) from its descriptor and the
// chunk origin `offs`.  Every coordinate is `(offs + c) as f32` exactly as the reference
// evaluates it (lib.rs:323-325; the run end is the sentinel/next cell, lib.rs:249,309).
MM_HD void expand_face_desc(const TabEntry *tab, long long ox, long long oy, long long oz, uint32_t d,
                            unsigned char *dst) {
	const uint32_t id = d & 0xffu;
	const int st = (int)((d >> 8) & 15u), c2 = (int)((d >> 12) & 15u), c1 = (int)((d >> 16) & 15u);
	const int p = (int)((d >> 20) & 7u), len = (int)((d >> 23) & 31u);
	const bool xin = (p == 1 || p == 4);
	const TabEntry e = tab[id];
	const long long oin = xin ? ox : oy, oout = xin ? oy : ox;
	const float last_v = i64_to_f32(oin + st);
	const float vlen = fsub(i64_to_f32(oin + st + len), last_v);  // vlen = v - last_v, lib.rs:249,256
	const float o1 = i64_to_f32(oout + c1), z = i64_to_f32(oz + c2);
	const float siz = 1.0f;  // lib.rs:314
	QuadArgs a;
	if (xin) {  // x-inner: row (y, z); run along x
		a.x = last_v; a.z = z;
		a.xsd = vlen; a.ysd = 0.0f; a.yd = 0.0f; a.zd = siz;
		a.tex = e.x & 0xffffu;                               // id_sides
		if (p == 1) { a.y = o1; a.rev = true; }              // lib.rs:342
		else { a.y = fadd(o1, siz); a.rev = false; }         // lib.rs:390
	} else {                 // y-inner: row (x, z); run along y
		a.y = last_v;
		if (p == 0) {        // lib.rs:326  (x, last_y, z), (siz, 0, ylen, 0), id_bottom
			a.x = o1; a.z = z; a.xsd = siz; a.ysd = 0.0f; a.yd = vlen; a.zd = 0.0f;
			a.tex = e.y & 0xffffu; a.rev = false;
		} else if (p == 2) { // lib.rs:358  (x, last_y, z), (0, ylen, 0, siz), id_sides
			a.x = o1; a.z = z; a.xsd = 0.0f; a.ysd = vlen; a.yd = 0.0f; a.zd = siz;
			a.tex = e.x & 0xffffu; a.rev = false;
		} else if (p == 3) { // lib.rs:374  (x, last_y, z + siz), (siz, 0, ylen, 0), id_top, rev
			a.x = o1; a.z = fadd(z, siz); a.xsd = siz; a.ysd = 0.0f; a.yd = vlen; a.zd = 0.0f;
			a.tex = e.x >> 16; a.rev = true;
		} else {             // lib.rs:406  (x + siz, last_y, z), (0, ylen, 0, siz), id_sides, rev
			a.x = fadd(o1, siz); a.z = z; a.xsd = 0.0f; a.ysd = vlen; a.yd = 0.0f; a.zd = siz;
			a.tex = e.x & 0xffffu; a.rev = true;
		}
	}
	a.us = fadd(a.xsd, a.ysd);  // tex_pos u = xsd + ysd, lib.rs:142
	a.vt = fadd(a.yd, a.zd);    // tex_pos v = yd + zd,  lib.rs:143
	write_quad(dst, a);
}

// Quad slot `part` (0..3) of a crossed plant: lib.rs:426-437.  Parts 0,1 are the
// forward and reverse halves of the first rpush_face_bidi!, parts 2,3 of the second.
MM_HD void expand_plant_desc(const TabEntry *tab, long long ox, long long oy, long long oz, uint32_t d, int part,
                             unsigned char *dst) {
	const TabEntry e = tab[d & 0xffu];
	const int cell = (int)((d >> 8) & 0xfffu);
	const int cx = cell >> 8, cy = (cell >> 4) & 15, cz = cell & 15;
	const float siz = 1.0f;
	const float sq = siz * 1.41421356237309504880168872420969808f * 0.5f;  // lib.rs:428 (exact: *1, *0.5)
	const float sqh = sq * 0.5f;                                           // lib.rs:429
	const float x = fadd(i64_to_f32(ox + cx), 0.5f * siz);                 // lib.rs:431
	const float y = fadd(i64_to_f32(oy + cy), 0.5f * siz);
	const float z = i64_to_f32(oz + cz);
	const float tsiz = siz * 0.5f;                                         // lib.rs:433
	const float h = siz * 0.95f;
	QuadArgs a;
	a.y = fsub(y, sqh); a.z = z;
	a.ysd = sq; a.yd = 0.0f; a.zd = h;
	if (part < 2) { a.x = fsub(x, sqh); a.xsd = sq; }    // lib.rs:435
	else { a.x = fadd(x, sqh); a.xsd = -sq; }            // lib.rs:437
	a.us = fadd(tsiz, tsiz);   // xstd + ystd, lib.rs:165
	a.vt = fadd(0.0f, h);      // ytd + ztd,   lib.rs:166
	a.tex = e.x & 0xffffu;
	a.rev = (part & 1) != 0;
	write_quad(dst, a);
}

}  // namespace mm
