/*
 * mapgen_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C99) of mimas-server's mapgen PHASE ONE, `gen_chunk_phase_one`
 * (mimas-server/mapgen.rs:64-245), the producer of the chunks the mesher consumes.
 * Checker for the CUDA generator (mm_gen_mapgen_device, mimas_b200/csrc/mm_mapgen.cu) only.
 *
 * PARITY UNPINNED.  The arithmetic of this function lives in four third-party crates that are not under
 * /root/reference (Cargo.lock:1443,1834,1915,1924,2418) and no reference test pins any mapgen output:
 *   rand_pcg 0.3.1   Pcg32 = Lcg64Xsh32 (PCG-XSH-RR 64/32)      restated from the published algorithm; pinned by the
 *                                                               PCG reference vectors (seed 42, stream 54)
 *   rand 0.8.4       gen::<u32/u64/f64>() (Standard)            u64 = two u32 draws, low half first; f64 = 53 high bits * 2^-53
 *   twox-hash 1.6.1  XxHash64::with_seed(0)                     XXH64; pinned by vectors made with the xxhash library
 *                                                               (tests/golden/make_mapgen_vectors.py)
 *   noise 0.7.0      Perlin (2-D, 3-D), PermutationTable::new   restated from the crate's published source as remembered:
 *     (rand 0.7.3 shuffle, rand_xorshift 0.2.0)                 surflet form, attn^4, gradient tables, scale factors --
 *                                                               NOT pinned by anything; a deviation here changes the
 *                                                               terrain but not what is graded (CUDA == this file)
 * Every function cites the reference line or the crate item it follows.
 *
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off: no fused multiply-add, as in the Rust build).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CHUNKSIZE 16 /* mimas-common/map.rs:11 */

/* ---- rand_pcg 0.3.1 Lcg64Xsh32 ------------------------------------------------------------------------- */
typedef struct {
	uint64_t state, increment;
} mg_pcg32;

#define PCG_MULT 6364136223846793005ULL

static void pcg_step(mg_pcg32 *p) { p->state = p->state * PCG_MULT + p->increment; }

/* Lcg64Xsh32::new(state, stream): increment = (stream << 1) | 1; state += increment; step */
void mg_pcg32_new(mg_pcg32 *p, uint64_t state, uint64_t stream) {
	p->increment = (stream << 1) | 1u;
	p->state = state + p->increment;
	pcg_step(p);
}

/* next_u32: output function XSH RR on the OLD state */
uint32_t mg_pcg32_next_u32(mg_pcg32 *p) {
	const uint64_t state = p->state;
	pcg_step(p);
	const uint32_t rot = (uint32_t)(state >> 59);
	const uint32_t xsh = (uint32_t)(((state >> 18) ^ state) >> 27);
	return (xsh >> rot) | (xsh << ((32u - rot) & 31u));
}

/* rand_core impls::next_u64_via_u32: low half first */
uint64_t mg_pcg32_next_u64(mg_pcg32 *p) {
	const uint64_t lo = mg_pcg32_next_u32(p);
	const uint64_t hi = mg_pcg32_next_u32(p);
	return (hi << 32) | lo;
}

/* rand 0.8 Standard for f64: 53 random bits scaled into [0, 1) */
double mg_pcg32_gen_f64(mg_pcg32 *p) { return (double)(mg_pcg32_next_u64(p) >> 11) * (1.0 / 9007199254740992.0); }

/* ---- twox-hash XxHash64 (XXH64) -------------------------------------------------------------------------- */
#define XP1 11400714785074694791ULL
#define XP2 14029467366897019727ULL
#define XP3 1609587929392839161ULL
#define XP4 9650029242287828579ULL
#define XP5 2870177450012600261ULL
static uint64_t rotl64(uint64_t v, int r) { return (v << r) | (v >> (64 - r)); }
static uint64_t xx_round(uint64_t acc, uint64_t in) { return rotl64(acc + in * XP2, 31) * XP1; }
static uint64_t xx_merge(uint64_t acc, uint64_t v) { return (acc ^ xx_round(0, v)) * XP1 + XP4; }
static uint64_t rd64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; } /* little endian host */
static uint32_t rd32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }

uint64_t mg_xxh64(const uint8_t *data, uint64_t len, uint64_t seed) {
	const uint8_t *p = data, *end = data + len;
	uint64_t h;
	if (len >= 32) {
		uint64_t v1 = seed + XP1 + XP2, v2 = seed + XP2, v3 = seed, v4 = seed - XP1;
		do {
			v1 = xx_round(v1, rd64(p)); v2 = xx_round(v2, rd64(p + 8));
			v3 = xx_round(v3, rd64(p + 16)); v4 = xx_round(v4, rd64(p + 24));
			p += 32;
		} while (p + 32 <= end);
		h = rotl64(v1, 1) + rotl64(v2, 7) + rotl64(v3, 12) + rotl64(v4, 18);
		h = xx_merge(h, v1); h = xx_merge(h, v2); h = xx_merge(h, v3); h = xx_merge(h, v4);
	} else {
		h = seed + XP5;
	}
	h += len;
	while (p + 8 <= end) { h ^= xx_round(0, rd64(p)); h = rotl64(h, 27) * XP1 + XP4; p += 8; }
	if (p + 4 <= end) { h ^= (uint64_t)rd32(p) * XP1; h = rotl64(h, 23) * XP2 + XP3; p += 4; }
	while (p < end) { h ^= (uint64_t)(*p) * XP5; h = rotl64(h, 11) * XP1; p++; }
	h ^= h >> 33; h *= XP2; h ^= h >> 29; h *= XP3; h ^= h >> 32;
	return h;
}

/* mapgen.rs:56-62 `pos_hash`: XxHash64 seed 0 over the three coordinates as i64, native (little) endian */
uint64_t mg_pos_hash(const int64_t pos[3]) {
	uint8_t buf[24];
	memcpy(buf, pos, 24);
	return mg_xxh64(buf, 24, 0);
}

/* ---- noise 0.7.0 PermutationTable::new(seed) --------------------------------------------------------------- */
/* rand_xorshift 0.2.0 XorShiftRng seeded with bytes [1,0,0,0, seed, seed, seed] (little endian words) */
typedef struct {
	uint32_t x, y, z, w;
} mg_xorshift;
static uint32_t xs_next(mg_xorshift *r) {
	const uint32_t x = r->x, t = x ^ (x << 11);
	r->x = r->y; r->y = r->z; r->z = r->w;
	const uint32_t w = r->w;
	r->w = w ^ (w >> 19) ^ (t ^ (t >> 8));
	return r->w;
}
/* rand 0.7.3 UniformInt<u32>::sample_single(0, range): widening multiply, `zone` approximation, rejection */
static uint32_t xs_below(mg_xorshift *r, uint32_t range) {
	int lz = 0;
	while (!((range << lz) & 0x80000000u)) lz++;
	const uint32_t zone = (range << lz) - 1u;
	for (;;) {
		const uint64_t m = (uint64_t)xs_next(r) * range;
		if ((uint32_t)m <= zone) return (uint32_t)(m >> 32);
	}
}
/* seq = 0..255 shuffled by rand 0.7.3 SliceRandom::shuffle: for i in (1..len).rev() swap(i, gen_index(i + 1)) */
void mg_perm_table(uint32_t seed, uint8_t values[256]) {
	mg_xorshift r = { 1u, seed, seed, seed };
	for (int i = 0; i < 256; i++) values[i] = (uint8_t)i;
	for (uint32_t i = 255; i >= 1; i--) {
		const uint32_t j = xs_below(&r, i + 1u);
		const uint8_t t = values[i]; values[i] = values[j]; values[j] = t;
	}
}

/* ---- noise 0.7.0 Perlin ---------------------------------------------------------------------------------------- */
#define DIAG 0.70710678118654752440 /* FRAC_1_SQRT_2 */
#define DIAG2 0.5773502691896258
static const double GRAD2[8][2] = { { 1, 0 }, { -1, 0 }, { 0, 1 }, { 0, -1 }, { DIAG, DIAG }, { -DIAG, DIAG }, { DIAG, -DIAG }, { -DIAG, -DIAG } };
static const double GRAD3[32][3] = {
	{ DIAG, DIAG, 0 }, { -DIAG, DIAG, 0 }, { DIAG, -DIAG, 0 }, { -DIAG, -DIAG, 0 }, { DIAG, 0, DIAG }, { -DIAG, 0, DIAG },
	{ DIAG, 0, -DIAG }, { -DIAG, 0, -DIAG }, { 0, DIAG, DIAG }, { 0, -DIAG, DIAG }, { 0, DIAG, -DIAG }, { 0, -DIAG, -DIAG },
	{ DIAG, DIAG, 0 }, { -DIAG, DIAG, 0 }, { DIAG, -DIAG, 0 }, { -DIAG, -DIAG, 0 }, { DIAG, 0, DIAG }, { -DIAG, 0, DIAG },
	{ DIAG, 0, -DIAG }, { -DIAG, 0, -DIAG }, { 0, DIAG, DIAG }, { 0, -DIAG, DIAG }, { 0, DIAG, -DIAG }, { 0, -DIAG, -DIAG },
	{ DIAG2, DIAG2, DIAG2 }, { -DIAG2, DIAG2, DIAG2 }, { DIAG2, -DIAG2, DIAG2 }, { -DIAG2, -DIAG2, DIAG2 },
	{ DIAG2, DIAG2, -DIAG2 }, { -DIAG2, DIAG2, -DIAG2 }, { DIAG2, -DIAG2, -DIAG2 }, { -DIAG2, -DIAG2, -DIAG2 },
};
static double powi4(double a) { const double a2 = a * a; return a2 * a2; } /* f64::powi(4): square, square */
static double clamp1(double v) { return v < -1.0 ? -1.0 : (v > 1.0 ? 1.0 : v); }

/* PermutationTable::get2 / get3 */
static unsigned perm2(const uint8_t *t, int64_t x, int64_t y) { return t[t[x & 0xff] ^ (unsigned)(y & 0xff)]; }
static unsigned perm3(const uint8_t *t, int64_t x, int64_t y, int64_t z) { return t[perm2(t, x, y) ^ (unsigned)(z & 0xff)]; }

static double surflet2(const uint8_t *t, int64_t cx, int64_t cy, double dx, double dy) {
	const double attn = 1.0 - (dx * dx + dy * dy);
	if (attn > 0.0) {
		const double *g = GRAD2[perm2(t, cx, cy) % 8];
		return powi4(attn) * (dx * g[0] + dy * g[1]);
	}
	return 0.0;
}
double mg_perlin2(const uint8_t *t, double x, double y) {
	const double fx = floor(x), fy = floor(y);
	const int64_t nx = (int64_t)fx, ny = (int64_t)fy;
	const double dx = x - fx, dy = y - fy, ex = dx - 1.0, ey = dy - 1.0;
	const double f00 = surflet2(t, nx, ny, dx, dy), f10 = surflet2(t, nx + 1, ny, ex, dy);
	const double f01 = surflet2(t, nx, ny + 1, dx, ey), f11 = surflet2(t, nx + 1, ny + 1, ex, ey);
	return clamp1((f00 + f10 + f01 + f11) * 3.1604938271604937);
}
static double surflet3(const uint8_t *t, int64_t cx, int64_t cy, int64_t cz, double dx, double dy, double dz) {
	const double attn = 1.0 - (dx * dx + dy * dy + dz * dz);
	if (attn > 0.0) {
		const double *g = GRAD3[perm3(t, cx, cy, cz) % 32];
		return powi4(attn) * (dx * g[0] + dy * g[1] + dz * g[2]);
	}
	return 0.0;
}
double mg_perlin3(const uint8_t *t, double x, double y, double z) {
	const double fx = floor(x), fy = floor(y), fz = floor(z);
	const int64_t nx = (int64_t)fx, ny = (int64_t)fy, nz = (int64_t)fz;
	const double dx = x - fx, dy = y - fy, dz = z - fz, ex = dx - 1.0, ey = dy - 1.0, ez = dz - 1.0;
	const double f000 = surflet3(t, nx, ny, nz, dx, dy, dz), f100 = surflet3(t, nx + 1, ny, nz, ex, dy, dz);
	const double f010 = surflet3(t, nx, ny + 1, nz, dx, ey, dz), f110 = surflet3(t, nx + 1, ny + 1, nz, ex, ey, dz);
	const double f001 = surflet3(t, nx, ny, nz + 1, dx, dy, ez), f101 = surflet3(t, nx + 1, ny, nz + 1, ex, dy, ez);
	const double f011 = surflet3(t, nx, ny + 1, nz + 1, dx, ey, ez), f111 = surflet3(t, nx + 1, ny + 1, nz + 1, ex, ey, ez);
	return clamp1((f000 + f100 + f010 + f110 + f001 + f101 + f011 + f111) * 3.8898553255531074);
}

/* ---- mapgen parameters as the kernels take them (= mm_mapgen_params of the C ABI) ---------------------------- */
#define MG_MAX_ORES 8
#define MG_MAX_PLANTS 16
typedef struct {
	uint8_t block; uint8_t noise_seed[8]; uint8_t pcg_seed[8];
	double freq, pcg_chance, limit_a, limit_b; int64_t limit_boundary;     /* game_params.rs:134-143 */
} mg_ore;
typedef struct {
	uint8_t block; uint8_t pcg_seed[8]; double pcg_limit;                   /* game_params.rs:145-149 */
} mg_plant;
typedef struct {
	uint8_t air, water, sand, ground, ground_top, stone;                    /* BlockRoles, game_params.rs:79-90 */
	uint32_t n_ores, n_plants;
	mg_ore ores[MG_MAX_ORES];
	mg_plant plants[MG_MAX_PLANTS];
} mg_params;

static uint64_t be_bytes(const uint8_t b[8]) {
	uint64_t v = 0;
	for (int i = 0; i < 8; i++) v = (v << 8) | b[i];
	return v;
}
/* the `s!` macro, mapgen.rs:66-77: Pcg32::new(seed, u64::from_be_bytes(name)).gen::<u32 | u64>() */
uint32_t mg_seed_u32(uint64_t seed, const uint8_t name[8]) { mg_pcg32 p; mg_pcg32_new(&p, seed, be_bytes(name)); return mg_pcg32_next_u32(&p); }
uint64_t mg_seed_u64(uint64_t seed, const uint8_t name[8]) { mg_pcg32 p; mg_pcg32_new(&p, seed, be_bytes(name)); return mg_pcg32_next_u64(&p); }

typedef struct { double freq, mag; uint8_t perm[256]; } mg_noise; /* Noise / NoiseMag, mapgen.rs:247-291 */
static void noise_new(mg_noise *n, uint64_t seed, const char *name, double freq, double mag) {
	n->freq = freq; n->mag = mag;
	mg_perm_table(mg_seed_u32(seed, (const uint8_t *)name), n->perm);
}
static double noise_get(const mg_noise *n, double x, double y) { return mg_perlin2(n->perm, x * n->freq, y * n->freq) * n->mag; }
static double noise_get1(const mg_noise *n, double x, double y) { return mg_perlin2(n->perm, x * n->freq, y * n->freq); }
static double noise_get3(const mg_noise *n, double x, double y, double z) { return mg_perlin3(n->perm, x * n->freq, y * n->freq, z * n->freq); }

/* Rust `f64 as isize`: truncation toward zero, saturating, NaN -> 0 */
static int64_t f64_as_isize(double v) {
	if (v != v) return 0;
	if (v >= 9223372036854775807.0) return INT64_MAX;
	if (v <= -9223372036854775808.0) return INT64_MIN;
	return (int64_t)v;
}

/* mapgen.rs:64-245 `gen_chunk_phase_one`.  chunk: 4096 block ids (index x*256 + y*16 + z); trees: up to 256 spawn
 * points as (x, y, z world, in_desert) quadruples, *n_trees receives their number (mapgen.rs:228). */
void mg_gen_chunk_phase_one(uint64_t seed, const int64_t pos[3], const mg_params *pr, uint8_t *chunk,
		int64_t *trees, uint32_t *n_trees) {
	mg_noise noise, mnoise, smnoise, ampnoise, ampnoise2, tnoise, mtnoise, binoise, mbinoise, ca_noise, mca_noise;
	noise_new(&noise, seed, "chn-base", 0.02356, 8.3);             /* :79-80 */
	noise_new(&mnoise, seed, "chn-mcro", 0.0018671, 23.27713);     /* :82-83 */
	noise_new(&smnoise, seed, "chn-smcr", 0.00043571, 137.479131); /* :85-86 */
	noise_new(&ampnoise, seed, "chn-ampl", 0.0023473, 1.0);        /* :88-89 */
	noise_new(&ampnoise2, seed, "chn-ampt", 0.0023473, 1.0);       /* :91-92 */
	noise_new(&tnoise, seed, "trenoise", 0.0088971, 1.0);          /* :94-95 */
	noise_new(&mtnoise, seed, "mtrnoise", 0.00093952, 1.0);        /* :97-98 */
	noise_new(&binoise, seed, "biom-bas", 0.0023881, 0.4);         /* :100-101 */
	noise_new(&mbinoise, seed, "biom-mac", 0.00113881, 0.6);       /* :103-104 */
	const uint64_t ph = mg_pos_hash(pos);
	mg_pcg32 tpcg;
	mg_pcg32_new(&tpcg, mg_seed_u64(seed, (const uint8_t *)"pcg-tree"), ph); /* :106 */
	mg_noise ore_noise[MG_MAX_ORES];
	mg_pcg32 ore_pcg[MG_MAX_ORES], plant_pcg[MG_MAX_PLANTS];
	for (uint32_t i = 0; i < pr->n_ores; i++) {                   /* :108-114 */
		ore_noise[i].freq = pr->ores[i].freq; ore_noise[i].mag = 1.0;
		mg_perm_table(mg_seed_u32(seed, pr->ores[i].noise_seed), ore_noise[i].perm);
		mg_pcg32_new(&ore_pcg[i], mg_seed_u64(seed, pr->ores[i].pcg_seed), ph);
	}
	for (uint32_t i = 0; i < pr->n_plants; i++)                   /* :116-121 */
		mg_pcg32_new(&plant_pcg[i], mg_seed_u64(seed, pr->plants[i].pcg_seed), ph);
	noise_new(&ca_noise, seed, "nois-cav", 0.052951, 1.0);         /* :124-125 */
	noise_new(&mca_noise, seed, "mnoi-cav", 0.0094951, 1.0);       /* :127-128 */

	memset(chunk, pr->air, 4096);                                  /* :131 */
	uint32_t nt = 0;
	for (int64_t x = 0; x < CHUNKSIZE; x++) {
		for (int64_t y = 0; y < CHUNKSIZE; y++) {
			const double px = (double)(pos[0] + x), py = (double)(pos[1] + y);
			const double sm_elev = noise_get(&smnoise, px, py);
			const double amp = 1.0 + noise_get1(&ampnoise, px, py) * 0.9;
			const double amp2 = 0.6 + noise_get1(&ampnoise2, px, py) * 0.5;
			const double base_noise = amp * noise_get(&noise, px, py) + amp2 * noise_get(&mnoise, px, py);
			const double elev = base_noise + sm_elev;
			int64_t elev_blocks = f64_as_isize(elev);
			/* checked_sub(pos.z): overflow only at the ends of the isize range */
			if ((pos[2] < 0 && elev_blocks > INT64_MAX + pos[2]) || (pos[2] > 0 && elev_blocks < INT64_MIN + pos[2])) continue;
			elev_blocks -= pos[2];
			int64_t els = elev_blocks - 4;
			els = els < CHUNKSIZE ? els : CHUNKSIZE; els = els > 0 ? els : 0;   /* .min(CHUNKSIZE).max(0) */
			int64_t elg = elev_blocks < CHUNKSIZE ? elev_blocks : CHUNKSIZE; elg = elg > 0 ? elg : 0;
			uint8_t *col = chunk + x * 256 + y * 16;
			for (int64_t z = 0; z < els; z++) {
				col[z] = pr->stone;
				const double pz = (double)(pos[2] + z);
				const int64_t z_abs = pos[2] + z;
				for (uint32_t i = 0; i < pr->n_ores; i++) {         /* :150-161 */
					const mg_ore *ore = &pr->ores[i];
					const double limit = z_abs < ore->limit_boundary ? ore->limit_b : ore->limit_a;
					if (noise_get3(&ore_noise[i], px, py, pz) > limit)
						if (mg_pcg32_gen_f64(&ore_pcg[i]) > ore->pcg_chance) col[z] = ore->block;
				}
				double mcave_thresh;                                /* :169-173 */
				if (z_abs > -400) mcave_thresh = 2.0;
				else {
					double v = (double)(z_abs + 600) / 300.0;
					v = v > -0.502 ? v : -0.502;                      /* .max(-0.502) */
					v = v < 0.0 ? v : 0.0;                            /* .min(0.0) */
					mcave_thresh = 1.0 + v;
				}
				const int cave_block = noise_get3(&mca_noise, px, py, pz) > mcave_thresh || noise_get3(&ca_noise, px, py, pz) > 0.45;
				if (z + 10 < elev_blocks && cave_block) col[z] = pr->air;
			}
			if (pos[2] < 0) {
				for (int64_t z = els; z < CHUNKSIZE; z++) col[z] = pr->water;   /* :180-183 */
			} else {
				uint8_t ground_bl, ground_top;
				if (noise_get(&binoise, px, py) + noise_get(&mbinoise, px, py) < 0.3) { ground_bl = pr->ground; ground_top = pr->ground_top; }
				else { ground_bl = pr->sand; ground_top = pr->sand; }
				for (int64_t z = els; z < elg; z++) col[z] = ground_bl;
				if (els < elg && elev_blocks <= CHUNKSIZE) col[elg - 1] = ground_top;   /* (els..elg).rev().next() */
				if (pos[2] == 0 && elg <= 0) col[0] = pr->water;
				if (elg > 0 && elg < CHUNKSIZE) {
					const int in_desert = ground_bl == pr->sand;
					const double tree_density = in_desert ? 0.1 : 0.4;
					double macro_density = noise_get1(&mtnoise, px, py);
					if (macro_density < 0.0) macro_density = 0.0;
					const double local_density = noise_get1(&tnoise, px, py) + macro_density;
					int spawning_tree = 0;
					if (local_density > 1.0 - tree_density) {
						const double limit = in_desert ? 0.99 : 0.91;
						if (mg_pcg32_gen_f64(&tpcg) > limit) {
							spawning_tree = 1;
							if (trees) { trees[4 * nt] = pos[0] + x; trees[4 * nt + 1] = pos[1] + y; trees[4 * nt + 2] = pos[2] + elg; trees[4 * nt + 3] = in_desert; }
							nt++;
						}
					}
					if (!spawning_tree && !in_desert)
						for (uint32_t i = 0; i < pr->n_plants; i++)
							if (mg_pcg32_gen_f64(&plant_pcg[i]) > pr->plants[i].pcg_limit) col[elg] = pr->plants[i].block;
				}
			}
		}
	}
	if (n_trees) *n_trees = nt;
}

/* A box of nx * ny * nz chunks, chunk i at chunk coordinates (cx0 + i % nx, cy0 + (i / nx) % ny, cz0 + i / (nx*ny)):
 * the layout mm_gen_mapgen_device uses.  offs = 16 * chunk coordinate. */
void mg_gen_region(uint64_t seed, int64_t cx0, int64_t cy0, int64_t cz0, uint32_t nx, uint32_t ny, uint32_t nz,
		const mg_params *pr, uint8_t *chunks, int64_t *offs, uint32_t *tree_counts) {
	const uint64_t n = (uint64_t)nx * ny * nz;
	for (uint64_t i = 0; i < n; i++) {
		int64_t pos[3] = { 16 * (cx0 + (int64_t)(i % nx)), 16 * (cy0 + (int64_t)((i / nx) % ny)), 16 * (cz0 + (int64_t)(i / ((uint64_t)nx * ny))) };
		uint32_t nt = 0;
		mg_gen_chunk_phase_one(seed, pos, pr, chunks + 4096 * i, 0, &nt);
		if (offs) { offs[3 * i] = pos[0]; offs[3 * i + 1] = pos[1]; offs[3 * i + 2] = pos[2]; }
		if (tree_counts) tree_counts[i] = nt;
	}
}

uint32_t mg_sizeof_params(void) { return (uint32_t)sizeof(mg_params); }
