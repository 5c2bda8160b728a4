// mm_mapgen.cu -- mimas-server mapgen PHASE ONE on the device: `gen_chunk_phase_one`
// (mimas-server/mapgen.rs:64-245) for a box of chunks, written straight into the mesher's input layout, so that
// server-side mapgen chunks reach the mesher without a host copy (the north star's batch submission path).
//
// What the reference computes sequentially -- per chunk 256 columns in (x, y) order, up to 16 cells each, with one
// PCG stream per ore, one for trees and one per plant that are drawn from only where a noise test passes -- is
// computed here by one CTA per chunk, thread <-> column: the noise tests of all cells first, then block-wide scans
// give every drawing cell its rank in its stream, and the stream is advanced to that rank in O(log rank)
// (an LCG jumps ahead by repeated squaring), so every thread draws exactly the numbers the sequential loop would.
//
// PARITY UNPINNED against the reference: the arithmetic lives in crates that are not under /root/reference
// (noise 0.7.0, rand_pcg 0.3.1, rand 0.8.4 / 0.7.3, rand_xorshift 0.2.0, twox-hash 1.6.1) and no reference test pins
// mapgen output.  PCG-XSH-RR 64/32 and XXH64 are pinned by their published vectors (tests/test_mapgen.py); the
// Perlin restatement (surflets, attn^4, gradient tables) is not.  Graded: bit-equal to oracle/mapgen_oracle.c.
// All floating point is f64 add / mul / div / floor / compare, never contracted (-fmad=false).
#include <cuda_runtime.h>
#include <math.h>
#include <string.h>

#include "../../include/mimas_meshgen_cuda.h"
#include "mm_internal.h"

namespace {

#define fail mm_fail
typedef unsigned long long u64;

constexpr int kMaxOres = MM_MAPGEN_MAX_ORES, kMaxPlants = MM_MAPGEN_MAX_PLANTS;
constexpr int kFixedNoises = 11;  // chn-base, chn-mcro, chn-smcr, chn-ampl, chn-ampt, trenoise, mtrnoise, biom-bas, biom-mac, nois-cav, mnoi-cav
enum { N_BASE, N_MACRO, N_SMACRO, N_AMP, N_AMP2, N_TREE, N_MTREE, N_BIOME, N_MBIOME, N_CAVE, N_MCAVE };

// ---- rand_pcg 0.3.1 Lcg64Xsh32 (PCG-XSH-RR 64/32) ------------------------------------------------------------
constexpr u64 kPcgMult = 6364136223846793005ull;
struct Pcg {
	u64 state, inc;
};
__host__ __device__ inline Pcg pcg_new(u64 state, u64 stream) {  // Lcg64Xsh32::new
	Pcg p;
	p.inc = (stream << 1) | 1ull;
	p.state = (state + p.inc) * kPcgMult + p.inc;
	return p;
}
__host__ __device__ inline uint32_t pcg_u32(Pcg &p) {
	const u64 s = p.state;
	p.state = s * kPcgMult + p.inc;
	const uint32_t rot = (uint32_t)(s >> 59), xsh = (uint32_t)(((s >> 18) ^ s) >> 27);
	return (xsh >> rot) | (xsh << ((32u - rot) & 31u));
}
__host__ __device__ inline u64 pcg_u64(Pcg &p) {  // rand_core next_u64_via_u32: low half first
	const u64 lo = pcg_u32(p), hi = pcg_u32(p);
	return (hi << 32) | lo;
}
__host__ __device__ inline double pcg_f64(Pcg &p) { return (double)(pcg_u64(p) >> 11) * (1.0 / 9007199254740992.0); }  // rand 0.8 Standard
// the stream `delta` steps further on: state' = A^delta * state + C * (A^delta - 1) / (A - 1), by repeated squaring
__device__ inline void pcg_advance(Pcg &p, u64 delta) {
	u64 acc_mult = 1, acc_plus = 0, cur_mult = kPcgMult, cur_plus = p.inc;
	while (delta) {
		if (delta & 1ull) {
			acc_mult *= cur_mult;
			acc_plus = acc_plus * cur_mult + cur_plus;
		}
		cur_plus = (cur_mult + 1ull) * cur_plus;
		cur_mult *= cur_mult;
		delta >>= 1;
	}
	p.state = acc_mult * p.state + acc_plus;
}

// ---- twox-hash XxHash64::with_seed(0) over three i64 (mapgen.rs:56-62 pos_hash): XXH64 of 24 bytes ----------
__host__ __device__ inline u64 rotl64(u64 v, int r) { return (v << r) | (v >> (64 - r)); }
__host__ __device__ inline u64 pos_hash(long long x, long long y, long long z) {
	const u64 P1 = 11400714785074694791ull, P2 = 14029467366897019727ull, P3 = 1609587929392839161ull,
	          P4 = 9650029242287828579ull, P5 = 2870177450012600261ull;
	u64 h = P5 + 24ull;  // seed 0, short input
	const u64 in[3] = {(u64)x, (u64)y, (u64)z};
	for (int i = 0; i < 3; i++) {
		h ^= rotl64(in[i] * P2, 31) * P1;
		h = rotl64(h, 27) * P1 + P4;
	}
	h ^= h >> 33; h *= P2; h ^= h >> 29; h *= P3; h ^= h >> 32;
	return h;
}

// ---- noise 0.7.0 ------------------------------------------------------------------------------------------------
// PermutationTable::new(seed): XorShiftRng seeded [1, seed, seed, seed], rand 0.7.3 shuffle of 0..255 (host, once per call)
struct XorShift {
	uint32_t x, y, z, w;
};
uint32_t xs_next(XorShift &r) {
	const uint32_t t = r.x ^ (r.x << 11);
	r.x = r.y; r.y = r.z; r.z = r.w;
	r.w = r.w ^ (r.w >> 19) ^ (t ^ (t >> 8));
	return r.w;
}
uint32_t xs_below(XorShift &r, uint32_t range) {  // UniformInt<u32>::sample_single(0, range)
	const uint32_t zone = (range << __builtin_clz(range)) - 1u;
	for (;;) {
		const u64 m = (u64)xs_next(r) * range;
		if ((uint32_t)m <= zone) return (uint32_t)(m >> 32);
	}
}
void perm_table(uint32_t seed, uint8_t v[256]) {
	XorShift r = {1u, seed, seed, seed};
	for (int i = 0; i < 256; i++) v[i] = (uint8_t)i;
	for (uint32_t i = 255; i >= 1; i--) {
		const uint32_t j = xs_below(r, i + 1u);
		const uint8_t t = v[i]; v[i] = v[j]; v[j] = t;
	}
}

// everything a chunk's generation needs besides its position; built on the host per call
struct GenTables {
	uint8_t perm[kFixedNoises + kMaxOres][256];
	double freq[kFixedNoises + kMaxOres], mag[kFixedNoises];
	u64 tree_state, ore_state[kMaxOres], plant_state[kMaxPlants];  // s!(name, u64): the streams' initial states
	double ore_chance[kMaxOres], ore_limit_a[kMaxOres], ore_limit_b[kMaxOres], plant_limit[kMaxPlants];
	long long ore_boundary[kMaxOres];
	uint8_t ore_block[kMaxOres], plant_block[kMaxPlants];
	uint8_t air, water, sand, ground, ground_top, stone;
	uint32_t n_ores, n_plants;
};

__constant__ double c_grad2[8][2];
__constant__ double c_grad3[32][3];

__device__ __forceinline__ double powi4(double a) {
	const double a2 = a * a;
	return a2 * a2;
}
__device__ __forceinline__ double clamp1(double v) { return v < -1.0 ? -1.0 : (v > 1.0 ? 1.0 : v); }
__device__ __forceinline__ unsigned perm2(const uint8_t *t, long long x, long long y) { return t[t[x & 0xff] ^ (unsigned)(y & 0xff)]; }

__device__ __forceinline__ double surflet2(const uint8_t *t, const double (*g2)[2], long long cx, long long cy, double dx, double dy) {
	const double attn = 1.0 - (dx * dx + dy * dy);
	if (attn > 0.0) {
		const unsigned g = perm2(t, cx, cy) & 7u;
		return powi4(attn) * (dx * g2[g][0] + dy * g2[g][1]);
	}
	return 0.0;
}
// g2 / g3: the gradient tables in SHARED memory (the index differs from lane to lane: from __constant__ memory
// such a read is served one address after the other)
__device__ double perlin2(const uint8_t *t, const double (*g2)[2], double x, double y) {
	const double fx = floor(x), fy = floor(y);
	const long long nx = (long long)fx, ny = (long long)fy;
	const double dx = x - fx, dy = y - fy, ex = dx - 1.0, ey = dy - 1.0;
	const double f00 = surflet2(t, g2, nx, ny, dx, dy), f10 = surflet2(t, g2, nx + 1, ny, ex, dy);
	const double f01 = surflet2(t, g2, nx, ny + 1, dx, ey), f11 = surflet2(t, g2, nx + 1, ny + 1, ex, ey);
	return clamp1((f00 + f10 + f01 + f11) * 3.1604938271604937);
}
// perlin3 (noise 0.7.0: eight surflets attn^4 * dot(gradient, d), attn = 1 - (dx*dx + dy*dy + dz*dz), gradient index
// t[t[t[x] ^ y] ^ z] & 31, sum * 3.8898553255531074, clamped; oracle/mapgen_oracle.c perlin3 is the literal form) for
// the cells of ONE column: x and y are fixed, so the floors, the fractions, the four (x, y) corner hashes and the
// partial sums dx*dx + dy*dy are worked out once per column and noise; every expression keeps the operand order of
// the literal form, so the values are the same bit for bit.
struct ColXY {
	double dx, dy, ex, ey, s00, s10, s01, s11;
	unsigned h00, h10, h01, h11;
};
__device__ __forceinline__ ColXY col_xy(const uint8_t *t, double x, double y) {
	const double fx = floor(x), fy = floor(y);
	const long long nx = (long long)fx, ny = (long long)fy;
	ColXY c;
	c.dx = x - fx; c.dy = y - fy; c.ex = c.dx - 1.0; c.ey = c.dy - 1.0;
	c.s00 = c.dx * c.dx + c.dy * c.dy; c.s10 = c.ex * c.ex + c.dy * c.dy;
	c.s01 = c.dx * c.dx + c.ey * c.ey; c.s11 = c.ex * c.ex + c.ey * c.ey;
	c.h00 = perm2(t, nx, ny); c.h10 = perm2(t, nx + 1, ny); c.h01 = perm2(t, nx, ny + 1); c.h11 = perm2(t, nx + 1, ny + 1);
	return c;
}
__device__ __forceinline__ double surflet3h(const uint8_t *t, const double (*g3)[3], unsigned h, long long cz, double dx, double dy,
                                            double dz, double sxy) {
	const double attn = 1.0 - (sxy + dz * dz);
	if (attn > 0.0) {
		const unsigned g = t[h ^ (unsigned)(cz & 0xff)] & 31u;
		return powi4(attn) * (dx * g3[g][0] + dy * g3[g][1] + dz * g3[g][2]);
	}
	return 0.0;
}
__device__ __forceinline__ double perlin3_col(const uint8_t *t, const double (*g3)[3], const ColXY &c, double z) {
	const double fz = floor(z);
	const long long nz = (long long)fz;
	const double dz = z - fz, ez = dz - 1.0;
	const double f000 = surflet3h(t, g3, c.h00, nz, c.dx, c.dy, dz, c.s00), f100 = surflet3h(t, g3, c.h10, nz, c.ex, c.dy, dz, c.s10);
	const double f010 = surflet3h(t, g3, c.h01, nz, c.dx, c.ey, dz, c.s01), f110 = surflet3h(t, g3, c.h11, nz, c.ex, c.ey, dz, c.s11);
	const double f001 = surflet3h(t, g3, c.h00, nz + 1, c.dx, c.dy, ez, c.s00), f101 = surflet3h(t, g3, c.h10, nz + 1, c.ex, c.dy, ez, c.s10);
	const double f011 = surflet3h(t, g3, c.h01, nz + 1, c.dx, c.ey, ez, c.s01), f111 = surflet3h(t, g3, c.h11, nz + 1, c.ex, c.ey, ez, c.s11);
	return clamp1((f000 + f100 + f010 + f110 + f001 + f101 + f011 + f111) * 3.8898553255531074);
}

__device__ __forceinline__ long long f64_as_isize(double v) {  // Rust `as isize`: toward zero, saturating, NaN -> 0
	if (v != v) return 0;
	if (v >= 9223372036854775807.0) return 0x7fffffffffffffffll;
	if (v <= -9223372036854775808.0) return (long long)0x8000000000000000ull;
	return (long long)v;
}

// exclusive scan of `v` over the 256 threads of the CTA in thread order; *total = the sum.  `sh`: 8 words of scratch.
__device__ uint32_t block_excl_scan(uint32_t v, uint32_t *sh, uint32_t *total) {
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
		if (lane >= d) inc += o;
	}
	__syncthreads();  // scratch free
	if (lane == 31) sh[warp] = inc;
	__syncthreads();
	uint32_t before = 0, all = 0;
#pragma unroll
	for (int w = 0; w < 8; w++) {
		const uint32_t t = sh[w];
		if (w < warp) before += t;
		all += t;
	}
	*total = all;
	return before + inc - v;
}

struct Box {
	u64 seed;
	long long cx0, cy0, cz0;
	uint32_t nx, ny, nz;
};

// one CTA of 256 threads per chunk, thread <-> column (x, y) = (tid >> 4, tid & 15): the reference's loop order
#ifndef MM_MAPGEN_MINBLOCKS
#define MM_MAPGEN_MINBLOCKS 5   // 48 registers, 5 CTAs per SM: measured best
#endif
__global__ void __launch_bounds__(256, MM_MAPGEN_MINBLOCKS) mapgen_kernel(const GenTables *__restrict__ gt, Box bx, uint32_t n, uint8_t *chunks,
                                                     long long *offs, uint32_t *tree_counts) {
	__shared__ uint8_t perm[kFixedNoises + kMaxOres][256];
	__shared__ uint32_t scratch[8];
	__shared__ double g2[8][2], g3[32][3];
	const int tid = threadIdx.x;
	if (tid < 16) (&g2[0][0])[tid] = (&c_grad2[0][0])[tid];
	if (tid < 96) (&g3[0][0])[tid] = (&c_grad3[0][0])[tid];
	const uint32_t n_ores = gt->n_ores, n_plants = gt->n_plants;
	for (uint32_t i = (uint32_t)tid; i < (kFixedNoises + n_ores) * 64u; i += 256u)
		reinterpret_cast<uint32_t *>(perm)[i] = reinterpret_cast<const uint32_t *>(gt->perm)[i];
	__syncthreads();
	const int x = tid >> 4, y = tid & 15;
	for (uint32_t ci = blockIdx.x; ci < n; ci += gridDim.x) {
		const long long ox = 16 * (bx.cx0 + (long long)(ci % bx.nx)), oy = 16 * (bx.cy0 + (long long)((ci / bx.nx) % bx.ny)),
		                oz = 16 * (bx.cz0 + (long long)(ci / (bx.nx * bx.ny)));
		if (tid < 3 && offs) offs[3 * (size_t)ci + tid] = tid == 0 ? ox : (tid == 1 ? oy : oz);
		const u64 ph = pos_hash(ox, oy, oz);
		const double px = (double)(ox + x), py = (double)(oy + y);
		auto n2 = [&](int k) { return perlin2(perm[k], g2, px * gt->freq[k], py * gt->freq[k]); };
		// mapgen.rs:136-143
		const double sm_elev = n2(N_SMACRO) * gt->mag[N_SMACRO];
		const double amp = 1.0 + n2(N_AMP) * 0.9;
		const double amp2 = 0.6 + n2(N_AMP2) * 0.5;
		const double base_noise = amp * (n2(N_BASE) * gt->mag[N_BASE]) + amp2 * (n2(N_MACRO) * gt->mag[N_MACRO]);
		const double elev = base_noise + sm_elev;
		long long elev_blocks = f64_as_isize(elev);
		const bool have = !((oz < 0 && elev_blocks > 0x7fffffffffffffffll + oz) || (oz > 0 && elev_blocks < (long long)0x8000000000000000ull + oz));
		elev_blocks -= have ? oz : 0;
		long long els = elev_blocks - 4;
		els = els < 16 ? els : 16; els = els > 0 ? els : 0;
		long long elg = elev_blocks < 16 ? elev_blocks : 16; elg = elg > 0 ? elg : 0;
		if (!have) els = elg = 0;
		uint8_t col[16];
#pragma unroll
		for (int z = 0; z < 16; z++) col[z] = gt->air;
		// underground: the noise tests of every cell first (mapgen.rs:146-178) ...
		uint32_t ore_pass[kMaxOres];
#pragma unroll
		for (int i = 0; i < kMaxOres; i++) ore_pass[i] = 0;
		uint32_t cave = 0;
		// (noise by noise, the column's cells inside: the tests are pure, their order is free)
#pragma unroll 1
		for (uint32_t i = 0; i < n_ores; i++) {
			const double f = gt->freq[kFixedNoises + i], la = gt->ore_limit_a[i], lb = gt->ore_limit_b[i];
			const long long boundary = gt->ore_boundary[i];
			const uint8_t *t = perm[kFixedNoises + i];
			const ColXY c = col_xy(t, px * f, py * f);
			uint32_t pass = 0;
			for (int z = 0; z < (int)els; z++) {
				const long long z_abs = oz + z;
				if (perlin3_col(t, g3, c, (double)z_abs * f) > (z_abs < boundary ? lb : la)) pass |= 1u << z;
			}
#pragma unroll
			for (int k = 0; k < kMaxOres; k++)
				if ((uint32_t)k == i) ore_pass[k] = pass;
		}
		{
			const double fm = gt->freq[N_MCAVE], fc = gt->freq[N_CAVE];
			const ColXY cm = col_xy(perm[N_MCAVE], px * fm, py * fm);
			uint32_t big = 0;
			for (int z = 0; z < (int)els; z++) {
				const long long z_abs = oz + z;
				double mcave_thresh;
				if (z_abs > -400) mcave_thresh = 2.0;
				else {
					double v = (double)(z_abs + 600) / 300.0;
					v = v > -0.502 ? v : -0.502;
					v = v < 0.0 ? v : 0.0;
					mcave_thresh = 1.0 + v;
				}
				if (perlin3_col(perm[N_MCAVE], g3, cm, (double)z_abs * fm) > mcave_thresh) big |= 1u << z;
			}
			const ColXY cc = col_xy(perm[N_CAVE], px * fc, py * fc);
			for (int z = 0; z < (int)els; z++) {
				const bool cave_block = ((big >> z) & 1u) || perlin3_col(perm[N_CAVE], g3, cc, (double)(oz + z) * fc) > 0.45;
				if (z + 10 < elev_blocks && cave_block) cave |= 1u << z;
			}
		}
		// ... then every ore stream is drawn from in (column, z) order: rank of this column's first draw by a block scan
#pragma unroll
		for (int z = 0; z < 16; z++)
			if (z < (int)els) col[z] = gt->stone;
		for (uint32_t i = 0; i < n_ores; i++) {
			uint32_t pass = 0;
#pragma unroll
			for (int k = 0; k < kMaxOres; k++)
				if ((uint32_t)k == i) pass = ore_pass[k];
			uint32_t total;
			const uint32_t before = block_excl_scan((uint32_t)__popc(pass), scratch, &total);
			if (pass) {
				Pcg p = pcg_new(gt->ore_state[i], ph);
				pcg_advance(p, 2ull * before);  // one f64 = two u32 draws
				const double chance = gt->ore_chance[i];
				const uint8_t blk = gt->ore_block[i];
#pragma unroll
				for (int z = 0; z < 16; z++)
					if ((pass >> z) & 1u)
						if (pcg_f64(p) > chance) col[z] = blk;
			}
		}
#pragma unroll
		for (int z = 0; z < 16; z++)
			if ((cave >> z) & 1u) col[z] = gt->air;
		// above the stone (mapgen.rs:179-243)
		bool tree_test = false, in_desert = false, surface = false;
		if (have) {
			if (oz < 0) {
#pragma unroll
				for (int z = 0; z < 16; z++)
					if (z >= (int)els) col[z] = gt->water;
			} else {
				const bool soil = n2(N_BIOME) * gt->mag[N_BIOME] + n2(N_MBIOME) * gt->mag[N_MBIOME] < 0.3;
				const uint8_t ground_bl = soil ? gt->ground : gt->sand, ground_top = soil ? gt->ground_top : gt->sand;
#pragma unroll
				for (int z = 0; z < 16; z++)
					if (z >= (int)els && z < (int)elg) col[z] = ground_bl;
				if (els < elg && elev_blocks <= 16) {
#pragma unroll
					for (int z = 0; z < 16; z++)
						if (z == (int)elg - 1) col[z] = ground_top;
				}
				if (oz == 0 && elg <= 0) col[0] = gt->water;
				if (elg > 0 && elg < 16) {
					surface = true;
					in_desert = ground_bl == gt->sand;
					const double tree_density = in_desert ? 0.1 : 0.4;
					double macro_density = n2(N_MTREE);
					if (macro_density < 0.0) macro_density = 0.0;
					const double local_density = n2(N_TREE) + macro_density;
					tree_test = local_density > 1.0 - tree_density;
				}
			}
		}
		// the tree stream, then the plant streams (drawn from only by the columns that did not get a tree)
		uint32_t total;
		uint32_t before = block_excl_scan(tree_test ? 1u : 0u, scratch, &total);
		bool spawning_tree = false;
		if (tree_test) {
			Pcg p = pcg_new(gt->tree_state, ph);
			pcg_advance(p, 2ull * before);
			spawning_tree = pcg_f64(p) > (in_desert ? 0.99 : 0.91);
		}
		uint32_t trees;
		block_excl_scan(spawning_tree ? 1u : 0u, scratch, &trees);
		const bool plants = surface && !spawning_tree && !in_desert;
		before = block_excl_scan(plants ? 1u : 0u, scratch, &total);
		if (plants) {
			uint8_t put = 0;
			bool any = false;
			for (uint32_t i = 0; i < n_plants; i++) {
				Pcg p = pcg_new(gt->plant_state[i], ph);
				pcg_advance(p, 2ull * before);
				if (pcg_f64(p) > gt->plant_limit[i]) {
					put = gt->plant_block[i];
					any = true;
				}
			}
			if (any) {
#pragma unroll
				for (int z = 0; z < 16; z++)
					if (z == (int)elg) col[z] = put;
			}
		}
		uint4 v;
		v.x = col[0] | (col[1] << 8) | (col[2] << 16) | ((uint32_t)col[3] << 24);
		v.y = col[4] | (col[5] << 8) | (col[6] << 16) | ((uint32_t)col[7] << 24);
		v.z = col[8] | (col[9] << 8) | (col[10] << 16) | ((uint32_t)col[11] << 24);
		v.w = col[12] | (col[13] << 8) | (col[14] << 16) | ((uint32_t)col[15] << 24);
		reinterpret_cast<uint4 *>(chunks + (size_t)ci * 4096u)[tid] = v;
		if (tid == 0 && tree_counts) tree_counts[ci] = trees;
	}
}

u64 be_bytes(const uint8_t b[8]) {
	u64 v = 0;
	for (int i = 0; i < 8; i++) v = (v << 8) | b[i];
	return v;
}
uint32_t seed_u32(u64 seed, const uint8_t name[8]) {  // the `s!` macro, mapgen.rs:66-77
	Pcg p = pcg_new(seed, be_bytes(name));
	return pcg_u32(p);
}
u64 seed_u64(u64 seed, const uint8_t name[8]) {
	Pcg p = pcg_new(seed, be_bytes(name));
	return pcg_u64(p);
}

}  // namespace

extern "C" {

int mm_gen_mapgen_device(mm_ctx *c, uint64_t seed, const mm_mapgen_params *pr, int64_t cx0, int64_t cy0, int64_t cz0, uint32_t nx,
                         uint32_t ny, uint32_t nz, uint8_t *d_chunks, int64_t *d_offs, uint32_t *d_tree_counts) {
	if (!c || !pr || !d_chunks || !nx || !ny || !nz) return MM_ERR_INVALID;
	if (pr->n_ores > MM_MAPGEN_MAX_ORES || pr->n_plants > MM_MAPGEN_MAX_PLANTS) return fail(c, MM_ERR_INVALID, "too many ores / plants");
	if ((uintptr_t)d_chunks & 15u) return fail(c, MM_ERR_INVALID, "chunks must be 16-byte aligned");
	MM_CUDA(c, cudaSetDevice(c->device));
	const uint64_t n64 = (uint64_t)nx * ny * nz;
	if (n64 > 0x7fffffffull) return fail(c, MM_ERR_INVALID, "region too large");
	static const char *kNames[kFixedNoises] = {"chn-base", "chn-mcro", "chn-smcr", "chn-ampl", "chn-ampt", "trenoise",
	                                           "mtrnoise", "biom-bas", "biom-mac", "nois-cav", "mnoi-cav"};
	static const double kFreq[kFixedNoises] = {0.02356, 0.0018671, 0.00043571, 0.0023473, 0.0023473, 0.0088971,
	                                           0.00093952, 0.0023881, 0.00113881, 0.052951, 0.0094951};  // mapgen.rs:79-128
	static const double kMag[kFixedNoises] = {8.3, 23.27713, 137.479131, 1.0, 1.0, 1.0, 1.0, 0.4, 0.6, 1.0, 1.0};
	GenTables *t = new GenTables();
	memset(t, 0, sizeof(*t));
	for (int k = 0; k < kFixedNoises; k++) {
		perm_table(seed_u32(seed, (const uint8_t *)kNames[k]), t->perm[k]);
		t->freq[k] = kFreq[k];
		t->mag[k] = kMag[k];
	}
	t->tree_state = seed_u64(seed, (const uint8_t *)"pcg-tree");
	for (uint32_t i = 0; i < pr->n_ores; i++) {
		const mm_mapgen_ore &o = pr->ores[i];
		perm_table(seed_u32(seed, o.noise_seed), t->perm[kFixedNoises + i]);
		t->freq[kFixedNoises + i] = o.freq;
		t->ore_state[i] = seed_u64(seed, o.pcg_seed);
		t->ore_chance[i] = o.pcg_chance; t->ore_limit_a[i] = o.limit_a; t->ore_limit_b[i] = o.limit_b;
		t->ore_boundary[i] = o.limit_boundary;
		t->ore_block[i] = o.block;
	}
	for (uint32_t i = 0; i < pr->n_plants; i++) {
		t->plant_state[i] = seed_u64(seed, pr->plants[i].pcg_seed);
		t->plant_limit[i] = pr->plants[i].pcg_limit;
		t->plant_block[i] = pr->plants[i].block;
	}
	t->air = pr->air; t->water = pr->water; t->sand = pr->sand; t->ground = pr->ground; t->ground_top = pr->ground_top; t->stone = pr->stone;
	t->n_ores = pr->n_ores; t->n_plants = pr->n_plants;
	static const double D = 0.70710678118654752440, D2 = 0.5773502691896258;  // gradient.rs of noise 0.7.0
	static const double g2[8][2] = {{1, 0}, {-1, 0}, {0, 1}, {0, -1}, {D, D}, {-D, D}, {D, -D}, {-D, -D}};
	double g3[32][3] = {{D, D, 0}, {-D, D, 0}, {D, -D, 0}, {-D, -D, 0}, {D, 0, D}, {-D, 0, D}, {D, 0, -D}, {-D, 0, -D},
	                    {0, D, D}, {0, -D, D}, {0, D, -D}, {0, -D, -D}};
	for (int i = 0; i < 12; i++)
		for (int k = 0; k < 3; k++) g3[12 + i][k] = g3[i][k];
	for (int i = 0; i < 8; i++) {
		g3[24 + i][0] = (i & 1) ? -D2 : D2;
		g3[24 + i][1] = (i & 2) ? -D2 : D2;
		g3[24 + i][2] = (i & 4) ? -D2 : D2;
	}
	int rc = MM_OK;
	GenTables *d_t = nullptr;
	auto body = [&]() -> int {
		MM_CUDA(c, cudaMemcpyToSymbolAsync(c_grad2, g2, sizeof(g2), 0, cudaMemcpyHostToDevice, c->stream));
		MM_CUDA(c, cudaMemcpyToSymbolAsync(c_grad3, g3, sizeof(g3), 0, cudaMemcpyHostToDevice, c->stream));
		MM_CUDA(c, cudaMalloc(&d_t, sizeof(GenTables)));
		MM_CUDA(c, cudaMemcpyAsync(d_t, t, sizeof(GenTables), cudaMemcpyHostToDevice, c->stream));
		const Box bx = {seed, cx0, cy0, cz0, nx, ny, nz};
		const uint32_t n = (uint32_t)n64;
		const uint32_t grid = n < (uint32_t)c->sm_count * 8u ? n : (uint32_t)c->sm_count * 8u;
		mapgen_kernel<<<grid, 256, 0, c->stream>>>(d_t, bx, n, d_chunks, (long long *)d_offs, d_tree_counts);
		MM_CUDA(c, cudaGetLastError());
		MM_CUDA(c, cudaStreamSynchronize(c->stream));  // the tables are freed below
		return MM_OK;
	};
	rc = body();
	cudaFree(d_t);
	delete t;
	return rc;
}

}  // extern "C"
