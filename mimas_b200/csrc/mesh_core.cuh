// mesh_core.cuh -- per-chunk data-path logic of the B200 chunk mesher.
//
// Everything in this header is written per *logical thread* (tid) between
// CTA-wide barriers, against an explicit shared-memory struct.  The CUDA kernel
// (mesh_kernel.cuh) calls these functions with tid = threadIdx.x and real
// barriers; tests/emul/ compiles the very same functions with g++ and runs the
// logical threads in a loop, so the bit-level logic can be checked on a machine
// without a GPU.  (That emulation is test infrastructure; the product library
// never runs it.)
//
// What is computed: mimas-meshgen/lib.rs:231-446 `mesh_for_chunk`
//   * six run-merging face passes (lib.rs:316-410) -> `intransparent`
//   * crossed-plant pass (lib.rs:412-440)          -> `transparent`
// re-formulated for a GPU as: per-row bit masks -> run-start masks -> CTA-wide
// exclusive scan in emission order -> 32-bit quad descriptors -> one thread per
// quad expands 216 bytes into a shared-memory staging tile that a bulk async
// copy (TMA) stores to the packed vertex stream.
#pragma once
#include <stdint.h>

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

constexpr int kThreads = 128;          // logical threads per CTA
constexpr int kRows = 256;             // rows per pass: (c1, c2) in 16 x 16
constexpr int kPasses = 6;             // lib.rs:316-410
constexpr int kTileQuads = 128;        // quads expanded per staging tile (one per thread)
constexpr int kQuadBytes = 216;        // 6 vertices x 36 B
constexpr int kQuadWords = 54;
constexpr int kDescCap = 1024;         // descriptors held in shared memory per round
constexpr int kPlantQuads = 4;         // 2 x rpush_face_bidi! = 24 vertices = 4 quad slots
constexpr int kStageBytes = kTileQuads * kQuadBytes;  // 27648, a multiple of 16

// Block-table entry as the kernel sees it (built from mm_block_entry on the host):
//   x = id_sides | id_top << 16   (for a crossed block: x = crossed texture id)
//   y = id_bottom | kind << 16    (kind: 0 none, 1 blocky, 2 crossed)
struct TabEntry {
	uint32_t x, y;
};

// Shared memory of one CTA.  Plain struct so that host emulation can allocate it.
struct alignas(128) Smem {
	// staging tile: +16 so the tile can be shifted by 8 B to match the 16-B phase of
	// its destination in the vertex stream (quads are only 8-B aligned: 216 = 8 * 27)
	alignas(128) unsigned char stage[kStageBytes + 16];
	alignas(16) unsigned char raw[4096];       // the chunk, index x*256 + y*16 + z (map.rs:83-86)
	alignas(16) TabEntry tab[256];             // block table
	alignas(16) uint32_t desc[kDescCap];       // quad / plant descriptors of the current round
	uint16_t S[kPasses][kRows];                // run-start mask per (pass, row), bit = inner coordinate
	uint16_t T[kPasses][kRows];                // run-boundary mask: start | not visible
	uint16_t pre[kPasses * kRows];             // counts, then exclusive prefix (emission order)
	uint16_t By[kRows];                        // blocky mask over y for row (x, z)
	uint16_t Bx[kRows];                        // blocky mask over x for row (y, z)
	uint16_t Cz[kRows];                        // crossed mask over z for column (x, y)
	uint16_t cpre[kRows];                      // plants: counts, then exclusive prefix (x, y order)
	float pos[3][20];                          // (offs + c) as f32 for c in 0..16, per axis
	uint32_t warp_sums[8];
	uint32_t n_quads, n_plants;                // chunk totals
	uint32_t any_crossed;
	unsigned long long base_q, base_t;         // first quad slot of this chunk in each stream
	unsigned long long ticket;
};

MM_HD int popc16(uint32_t v) {
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

// ---- phase 1: world coordinates --------------------------------------------------
// pos[a][c] = (offs[a] + c) as f32 for c in 0..=16: every `pos.x as f32` the reference
// evaluates (lib.rs:323-325 etc.), including the sentinel cell 16 (lib.rs:309).
MM_HD void phase_pos(Smem &s, int tid, const long long *offs3) {
	if (tid < 51) {
		int a = tid / 17, c = tid % 17;
		s.pos[a][c] = i64_to_f32(offs3[a] + (long long)c);
	}
}

// ---- phase 2: per-row masks --------------------------------------------------------
// Registers a logical thread carries from phase 2 to phase 3 for one row index r.
struct RowRegs {
	uint32_t Es_Et;  // low 16: sides-texture equality along y, high 16: top-texture
	uint32_t Eb_Ex;  // low 16: bottom-texture equality along y, high 16: sides equality along x
	uint32_t By_Bx;  // low 16: blocky over y (row (x,z)), high 16: blocky over x (row (y,z))
};

// Row r is (c1, c2) = (r >> 4, r & 15).  For the y-inner passes (1,3,4,6; lib.rs:318)
// that is (x, z); for the x-inner passes (2,5; lib.rs:334) it is (y, z).
// Bit i of an equality mask says: cell i and cell i-1 of the row have the same
// texture id for that face (bit 0 is never used).
MM_HD RowRegs phase_rows(Smem &s, int r, uint32_t &any_crossed) {
	const int c1 = r >> 4, c2 = r & 15;
	RowRegs o;
	uint32_t By = 0, Es = 0, Et = 0, Eb = 0, cr = 0;
	TabEntry prev = {0xffffffffu, 0xffffffffu};
#pragma unroll
	for (int y = 0; y < 16; y++) {  // row (x = c1, z = c2), cells along y
		const TabEntry e = s.tab[s.raw[c1 * 256 + y * 16 + c2]];
		const uint32_t kind = e.y >> 16;
		By |= (kind == 1u ? 1u : 0u) << y;
		cr |= (kind == 2u ? 1u : 0u);
		Es |= (((e.x ^ prev.x) & 0xffffu) == 0u ? 1u : 0u) << y;
		Et |= (((e.x ^ prev.x) >> 16) == 0u ? 1u : 0u) << y;
		Eb |= (((e.y ^ prev.y) & 0xffffu) == 0u ? 1u : 0u) << y;
		prev = e;
	}
	uint32_t Bx = 0, Ex = 0;
	prev.x = 0xffffffffu;
#pragma unroll
	for (int x = 0; x < 16; x++) {  // row (y = c1, z = c2), cells along x
		const TabEntry e = s.tab[s.raw[x * 256 + c1 * 16 + c2]];
		Bx |= ((e.y >> 16) == 1u ? 1u : 0u) << x;
		Ex |= (((e.x ^ prev.x) & 0xffffu) == 0u ? 1u : 0u) << x;
		prev = e;
	}
	s.By[r] = (uint16_t)By;
	s.Bx[r] = (uint16_t)Bx;
	any_crossed |= cr;
	o.Es_Et = Es | (Et << 16);
	o.Eb_Ex = Eb | (Ex << 16);
	o.By_Bx = By | (Bx << 16);
	return o;
}

// ---- phase 3: visibility, run starts, counts ---------------------------------------
// lib.rs:273-293: a face is visible iff the cell is blocky and the neighbour in the
// pass direction is outside the chunk or not blocky.  A run (lib.rs:244-271) starts
// at cell i iff it is visible and (cell i-1 is not visible or has another texture).
MM_HD void start_masks(Smem &s, int p, int r, uint32_t V, uint32_t E) {
	const uint32_t S = V & ~((V << 1) & E);
	const uint32_t T = (S | ~V) & 0xffffu;
	s.S[p][r] = (uint16_t)S;
	s.T[p][r] = (uint16_t)T;
	s.pre[p * kRows + r] = (uint16_t)popc16(S);
}

MM_HD void phase_vis(Smem &s, int r, const RowRegs &g) {
	const int c1 = r >> 4, c2 = r & 15;
	const uint32_t By = g.By_Bx & 0xffffu, Bx = g.By_Bx >> 16;
	// y-inner rows: (x, z) = (c1, c2)
	const uint32_t Bzm = c2 > 0 ? s.By[r - 1] : 0u;    // (x, *, z-1)
	const uint32_t Bzp = c2 < 15 ? s.By[r + 1] : 0u;   // (x, *, z+1)
	const uint32_t Bxm = c1 > 0 ? s.By[r - 16] : 0u;   // (x-1, *, z)
	const uint32_t Bxp = c1 < 15 ? s.By[r + 16] : 0u;  // (x+1, *, z)
	start_masks(s, 0, r, By & ~Bzm, g.Eb_Ex & 0xffffu);  // pass 1, [0,0,-1], id_bottom (lib.rs:317-330)
	start_masks(s, 2, r, By & ~Bxm, g.Es_Et & 0xffffu);  // pass 3, [-1,0,0], id_sides  (lib.rs:349-362)
	start_masks(s, 3, r, By & ~Bzp, g.Es_Et >> 16);      // pass 4, [0,0,1],  id_top    (lib.rs:365-378)
	start_masks(s, 5, r, By & ~Bxp, g.Es_Et & 0xffffu);  // pass 6, [1,0,0],  id_sides  (lib.rs:397-410)
	// x-inner rows: (y, z) = (c1, c2)
	const uint32_t Bym = c1 > 0 ? s.Bx[r - 16] : 0u;   // (*, y-1, z)
	const uint32_t Byp = c1 < 15 ? s.Bx[r + 16] : 0u;  // (*, y+1, z)
	start_masks(s, 1, r, Bx & ~Bym, g.Eb_Ex >> 16);      // pass 2, [0,-1,0], id_sides  (lib.rs:333-346)
	start_masks(s, 4, r, Bx & ~Byp, g.Eb_Ex >> 16);      // pass 5, [0,1,0],  id_sides  (lib.rs:381-394)
}

// Crossed plants (lib.rs:414-425): column (x, y) = (col >> 4, col & 15), bit z.
MM_HD void phase_crossed_cols(Smem &s, int col) {
	uint32_t C = 0;
#pragma unroll
	for (int z = 0; z < 16; z++)
		C |= ((s.tab[s.raw[col * 16 + z]].y >> 16) == 2u ? 1u : 0u) << z;
	s.Cz[col] = (uint16_t)C;
	s.cpre[col] = (uint16_t)popc16(C);
}

// ---- descriptors -------------------------------------------------------------------
// quad:  pass(3) | row(8) << 3 | start(4) << 11 | (len-1)(4) << 15 | block id(8) << 19
// plant: cell index(12) | block id(8) << 12
MM_HD void phase_desc_row(Smem &s, int r, uint32_t w0) {
	const int c1 = r >> 4, c2 = r & 15;
#pragma unroll 1
	for (int p = 0; p < kPasses; p++) {
		uint32_t S = s.S[p][r];
		if (S == 0u) continue;
		uint32_t slot = s.pre[p * kRows + r];
		if (slot >= w0 + (uint32_t)kDescCap || slot + (uint32_t)popc16(S) <= w0) continue;
		const uint32_t T = s.T[p][r];
		const bool xin = (p == 1 || p == 4);
		while (S) {
			const int st = ffs32(S) - 1;
			S &= S - 1u;
			if (slot >= w0 && slot < w0 + (uint32_t)kDescCap) {
				const uint32_t m = T >> (st + 1);
				const int len = m ? ffs32(m) : 16 - st;
				const int cell = xin ? (st * 256 + c1 * 16 + c2) : (c1 * 256 + st * 16 + c2);
				const uint32_t id = s.raw[cell];
				s.desc[slot - w0] = (uint32_t)p | ((uint32_t)r << 3) | ((uint32_t)st << 11) |
				                    ((uint32_t)(len - 1) << 15) | (id << 19);
			}
			slot++;
		}
	}
}

MM_HD void phase_desc_col(Smem &s, int col, uint32_t w0) {
	uint32_t C = s.Cz[col];
	uint32_t slot = s.cpre[col];
	while (C) {
		const int z = ffs32(C) - 1;
		C &= C - 1u;
		if (slot >= w0 && slot < w0 + (uint32_t)kDescCap) {
			const uint32_t cell = (uint32_t)col * 16u + (uint32_t)z;
			s.desc[slot - w0] = cell | ((uint32_t)s.raw[cell] << 12);
		}
		slot++;
	}
}

// ---- quad expansion -----------------------------------------------------------------
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

// One quad of the six face passes from its descriptor (pass table: SURVEY.md §8a A5).
MM_HD void expand_face_quad(const Smem &s, uint32_t d, unsigned char *dst) {
	const int p = d & 7, r = (d >> 3) & 255, st = (d >> 11) & 15, len = ((d >> 15) & 15) + 1;
	const TabEntry e = s.tab[(d >> 19) & 255];
	const int c1 = r >> 4, c2 = r & 15;
	const float siz = 1.0f;  // lib.rs:314
	QuadArgs a;
	if (p == 1 || p == 4) {  // x-inner: row (y, z); run along x
		const float last_x = s.pos[0][st];
		const float xlen = fsub(s.pos[0][st + len], last_x);  // vlen = v - last_v, lib.rs:249,256
		const float y = s.pos[1][c1], z = s.pos[2][c2];
		a.x = last_x; a.z = z;
		a.xsd = xlen; a.ysd = 0.0f; a.yd = 0.0f; a.zd = siz;
		a.tex = e.x & 0xffffu;                               // id_sides
		if (p == 1) { a.y = y; a.rev = true; }               // lib.rs:342
		else { a.y = fadd(y, siz); a.rev = false; }          // lib.rs:390
	} else {                 // y-inner: row (x, z); run along y
		const float last_y = s.pos[1][st];
		const float ylen = fsub(s.pos[1][st + len], last_y);
		const float x = s.pos[0][c1], z = s.pos[2][c2];
		a.y = last_y;
		if (p == 0) {        // lib.rs:326  (x, last_y, z), (siz, 0, ylen, 0), id_bottom
			a.x = x; a.z = z; a.xsd = siz; a.ysd = 0.0f; a.yd = ylen; a.zd = 0.0f;
			a.tex = e.y & 0xffffu; a.rev = false;
		} else if (p == 2) { // lib.rs:358  (x, last_y, z), (0, ylen, 0, siz), id_sides
			a.x = x; a.z = z; a.xsd = 0.0f; a.ysd = ylen; a.yd = 0.0f; a.zd = siz;
			a.tex = e.x & 0xffffu; a.rev = false;
		} else if (p == 3) { // lib.rs:374  (x, last_y, z + siz), (siz, 0, ylen, 0), id_top, rev
			a.x = x; a.z = fadd(z, siz); a.xsd = siz; a.ysd = 0.0f; a.yd = ylen; a.zd = 0.0f;
			a.tex = e.x >> 16; a.rev = true;
		} else {             // lib.rs:406  (x + siz, last_y, z), (0, ylen, 0, siz), id_sides, rev
			a.x = fadd(x, siz); a.z = z; a.xsd = 0.0f; a.ysd = ylen; a.yd = 0.0f; a.zd = siz;
			a.tex = e.x & 0xffffu; a.rev = true;
		}
	}
	a.us = fadd(a.xsd, a.ysd);  // tex_pos u = xsd + ysd, lib.rs:142
	a.vt = fadd(a.yd, a.zd);    // tex_pos v = yd + zd,  lib.rs:143
	write_quad(dst, a);
}

// Quad slot `part` (0..3) of a crossed plant: lib.rs:426-437.  Parts 0,1 are the
// forward and reverse halves of the first rpush_face_bidi! (X-Z), parts 2,3 of the
// second (Y-Z).
MM_HD void expand_plant_quad(const Smem &s, uint32_t d, int part, unsigned char *dst) {
	const int cell = d & 4095;
	const TabEntry e = s.tab[(d >> 12) & 255];
	const int cx = cell >> 8, cy = (cell >> 4) & 15, cz = cell & 15;
	const float siz = 1.0f;
	const float sq = siz * 1.41421356237309504880168872420969808f * 0.5f;  // lib.rs:428 (exact: *1, *0.5)
	const float sqh = sq * 0.5f;                                           // lib.rs:429
	const float x = fadd(s.pos[0][cx], 0.5f * siz);                        // lib.rs:431
	const float y = fadd(s.pos[1][cy], 0.5f * siz);
	const float z = s.pos[2][cz];
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
