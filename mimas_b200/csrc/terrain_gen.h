// terrain_gen.h -- synthetic "mapgen look-alike" chunk generator (benchmark input).
//
// NOT a replica of mimas-server's mapgen (mimas-server/mapgen.rs:64-245): that one
// draws from the `noise`, `rand_pcg`, `rand` and `twox-hash` crates, none of which are
// available here, and no reference test pins its output.  This generator only mimics
// its structure so that the mesher sees realistic chunks:
//   * per-column elevation from three octaves of 2-D value noise (cf. mapgen.rs:134-141)
//   * stone below elev-4 with ore veins by 3-D noise threshold + depth limits
//     (mapgen.rs:146-160, game-params.toml [[mapgen.ore]])
//   * caves carved where z + 10 < elev (mapgen.rs:169-172)
//   * chunks below z = 0 are flooded above the stone (mapgen.rs:175-178)
//   * ground / sand by a biome noise, grass on top (mapgen.rs:180-191)
//   * crossed plants on the surface (mapgen.rs:226-232) and small trees / cacti
//     (mimas-common/schematic.rs:24-45), kept inside the chunk
// Integer arithmetic only, so the host (gcc) and device (nvcc) builds of these
// functions produce identical bytes.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TG_HD __host__ __device__ __forceinline__
#else
#define TG_HD static inline
#endif

namespace tg {

// block ids of the default game (mimas-common/game_params.rs:240-252)
enum : uint8_t {
	AIR = 0, WATER = 1, SAND = 2, GROUND = 3, STONE = 5, LEAVES = 6, TREE = 7, CACTUS = 8, COAL = 9,
	IRON = 10, GRASSY = 11, COPPER = 13, TIN = 14, GOLD = 15, DIAMOND = 16, GRASS_1 = 43, FLOWER_1 = 48
};

TG_HD uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	return z ^ (z >> 31);
}
TG_HD uint32_t hash3(uint64_t seed, uint32_t salt, int64_t x, int64_t y, int64_t z) {
	uint64_t h = mix64(seed ^ (0x9e3779b97f4a7c15ull * (uint64_t)(salt + 1u)));
	h = mix64(h ^ (uint64_t)x);
	h = mix64(h ^ ((uint64_t)y * 0xc2b2ae3d27d4eb4full));
	h = mix64(h ^ ((uint64_t)z * 0x165667b19e3779f9ull));
	return (uint32_t)(h >> 32);
}
// floor division / modulo by a power of two for signed coordinates
TG_HD int64_t fdiv(int64_t v, int sh) { return v >= 0 ? (v >> sh) : -(((-v) + ((int64_t)1 << sh) - 1) >> sh); }

// smoothstep weight in 16-bit fixed point for a fraction f / 2^sh
TG_HD uint32_t smooth16(uint32_t f, int sh) {
	const uint64_t t = ((uint64_t)f << 16) >> sh;                 // 0 .. 65535
	return (uint32_t)((t * t * (3ull * 65536ull - 2ull * t)) >> 32);  // 0 .. 65536
}
TG_HD uint32_t lerp16(uint32_t a, uint32_t b, uint32_t w) {       // a,b in 0..65535; w in 0..65536
	return (uint32_t)(((uint64_t)a * (65536u - w) + (uint64_t)b * w) >> 16);
}
// 2-D value noise with lattice period 2^sh, result 0..65535
TG_HD uint32_t vnoise2(uint64_t seed, uint32_t salt, int64_t x, int64_t y, int sh) {
	const int64_t cx = fdiv(x, sh), cy = fdiv(y, sh);
	const uint32_t fx = (uint32_t)(x - (cx << sh)), fy = (uint32_t)(y - (cy << sh));
	const uint32_t wx = smooth16(fx, sh), wy = smooth16(fy, sh);
	const uint32_t v00 = hash3(seed, salt, cx, cy, 0) & 0xffffu, v10 = hash3(seed, salt, cx + 1, cy, 0) & 0xffffu;
	const uint32_t v01 = hash3(seed, salt, cx, cy + 1, 0) & 0xffffu, v11 = hash3(seed, salt, cx + 1, cy + 1, 0) & 0xffffu;
	return lerp16(lerp16(v00, v10, wx), lerp16(v01, v11, wx), wy);
}
// 3-D value noise with lattice period 2^sh, result 0..65535
TG_HD uint32_t vnoise3(uint64_t seed, uint32_t salt, int64_t x, int64_t y, int64_t z, int sh) {
	const int64_t cx = fdiv(x, sh), cy = fdiv(y, sh), cz = fdiv(z, sh);
	const uint32_t wx = smooth16((uint32_t)(x - (cx << sh)), sh), wy = smooth16((uint32_t)(y - (cy << sh)), sh),
	               wz = smooth16((uint32_t)(z - (cz << sh)), sh);
	uint32_t v[2];
	for (int k = 0; k < 2; k++) {
		const uint32_t a = hash3(seed, salt, cx, cy, cz + k) & 0xffffu, b = hash3(seed, salt, cx + 1, cy, cz + k) & 0xffffu;
		const uint32_t c = hash3(seed, salt, cx, cy + 1, cz + k) & 0xffffu, d = hash3(seed, salt, cx + 1, cy + 1, cz + k) & 0xffffu;
		v[k] = lerp16(lerp16(a, b, wx), lerp16(c, d, wx), wy);
	}
	return lerp16(v[0], v[1], wz);
}
// amplitude * (noise - 0.5), in blocks
TG_HD int64_t centred(uint32_t n, int64_t amp2) { return (int64_t)(((uint64_t)n * (uint64_t)amp2) >> 16) - amp2 / 2; }

struct Column {
	int64_t elev;   // world z of the first air block above ground (mapgen.rs `elev_blocks`)
	bool desert;
	uint8_t plant;  // 0 or a crossed plant id to put on the surface
	uint8_t tree;   // 0 none, 1 tree (trunk 7 + leaves 6), 2 cactus (8)
};

TG_HD Column column_info(uint64_t seed, int64_t wx, int64_t wy) {
	Column c;
	c.elev = -4 + centred(vnoise2(seed, 1, wx, wy, 8), 56) + centred(vnoise2(seed, 2, wx, wy, 6), 28) +
	         centred(vnoise2(seed, 3, wx, wy, 4), 9);
	c.desert = vnoise2(seed, 4, wx, wy, 7) > 42000u;
	const uint32_t h = hash3(seed, 5, wx, wy, 0);
	c.plant = 0;
	c.tree = 0;
	if (!c.desert) {
		if ((h & 0xffffu) < 650u) c.tree = 1;                                    // ~1 % of columns
		else if (((h >> 16) & 0xffffu) < 2600u) c.plant = (uint8_t)(GRASS_1 + (h % 13u));  // ~4 %
	} else if ((h & 0xffffu) < 200u) {
		c.tree = 2;
	}
	return c;
}

TG_HD uint8_t underground_block(uint64_t seed, int64_t wx, int64_t wy, int64_t wz, int64_t elev) {
	// caves (mapgen.rs:161-172): only with 10 blocks of cover
	if (wz + 10 < elev && vnoise3(seed, 20, wx, wy, wz, 4) > 47500u) return AIR;
	// ore veins: blob noise above a limit and a per-cell coin (mapgen.rs:149-159)
	const uint32_t vein = vnoise3(seed, 21, wx, wy, wz, 3);
	if (vein > 45000u) {
		const uint32_t h = hash3(seed, 22, wx, wy, wz);
		const uint32_t kind = hash3(seed, 23, fdiv(wx, 3), fdiv(wy, 3), fdiv(wz, 3)) % 100u;
		if ((h & 0xffu) < 150u) {
			if (kind < 40u) return COAL;
			if (kind < 65u) return IRON;
			if (kind < 80u) return COPPER;
			if (kind < 92u) return TIN;
			if (wz < -16) return kind < 97u ? GOLD : DIAMOND;
			return COAL;
		}
	}
	return STONE;
}

// One column of one chunk: 16 block ids for z = 0..15 at chunk origin (ox, oy, oz) + (x, y).
// Follows the per-column control flow of mapgen.rs:134-243.
TG_HD void gen_column(uint64_t seed, int64_t ox, int64_t oy, int64_t oz, int x, int y, const Column &ci,
                      uint8_t *col /*[16]*/) {
	const int64_t wx = ox + x, wy = oy + y;
	const int64_t e = ci.elev - oz;  // elev_blocks relative to the chunk
	int64_t els = e - 4;
	els = els < 0 ? 0 : (els > 16 ? 16 : els);
	const int64_t elg = e < 0 ? 0 : (e > 16 ? 16 : e);
	for (int z = 0; z < 16; z++) col[z] = AIR;
	for (int z = 0; z < (int)els; z++) col[z] = underground_block(seed, wx, wy, oz + z, ci.elev);
	if (oz < 0) {
		for (int z = (int)els; z < 16; z++) col[z] = WATER;
	} else {
		const uint8_t ground = ci.desert ? SAND : GROUND, top = ci.desert ? SAND : GRASSY;
		for (int z = (int)els; z < (int)elg; z++) col[z] = ground;
		if (elg > els && e <= 16) col[elg - 1] = top;
		if (oz == 0 && elg <= 0) col[0] = WATER;
		if (elg > 0 && elg < 16 && ci.plant) col[elg] = ci.plant;
	}
}

// Trees are written after all columns of the chunk exist: leaves first (into air only),
// then trunks, so the result does not depend on the order columns are visited in.
// Returns true if column (x, y) carries a tree that fits inside this chunk.
TG_HD bool tree_fits(int64_t oz, int x, int y, const Column &ci, int *base_z) {
	if (ci.tree == 0 || oz < 0) return false;
	const int64_t e = ci.elev - oz;
	if (e <= 0 || e > 9) return false;
	if (x < 2 || x > 13 || y < 2 || y > 13) return false;
	*base_z = (int)e;
	return true;
}

struct Region {
	uint64_t seed;
	long long cx0, cy0, cz0;
	uint32_t nx, ny, nz;
};

// leaves (into air only), then trunk / cactus: shared by the host loop and the device kernel
TG_HD void tree_leaves(uint8_t *chunk, int x, int y, int base, const Column &ci) {
	if (ci.tree != 1) return;
	for (int dz = 3; dz <= 5; dz++) {
		const int r = dz < 5 ? 1 : 0;
		for (int dx = -r; dx <= r; dx++)
			for (int dy = -r; dy <= r; dy++) {
				uint8_t *p = &chunk[(x + dx) * 256 + (y + dy) * 16 + base + dz];
				if (*p == AIR) *p = LEAVES;
			}
	}
}
TG_HD void tree_trunk(uint8_t *chunk, int x, int y, int base, const Column &ci) {
	const int h = ci.tree == 1 ? 4 : 3;
	for (int dz = 0; dz < h; dz++) chunk[x * 256 + y * 16 + base + dz] = ci.tree == 1 ? TREE : CACTUS;
}

// Chunk i of a region on the host (the device kernel in mm_extra.cu visits the same columns with one thread each).
static inline void gen_chunk_host(const Region &rg, uint32_t i, uint8_t *chunk, long long *offs) {
	const long long cx = rg.cx0 + (long long)(i % rg.nx), cy = rg.cy0 + (long long)((i / rg.nx) % rg.ny),
	                cz = rg.cz0 + (long long)(i / (rg.nx * rg.ny));
	const long long ox = cx * 16, oy = cy * 16, oz = cz * 16;
	offs[0] = ox; offs[1] = oy; offs[2] = oz;
	Column cols[256];
	for (int x = 0; x < 16; x++)
		for (int y = 0; y < 16; y++) {
			cols[x * 16 + y] = column_info(rg.seed, ox + x, oy + y);
			gen_column(rg.seed, ox, oy, oz, x, y, cols[x * 16 + y], chunk + x * 256 + y * 16);
		}
	int base;
	for (int k = 0; k < 256; k++)
		if (tree_fits(oz, k >> 4, k & 15, cols[k], &base)) tree_leaves(chunk, k >> 4, k & 15, base, cols[k]);
	for (int k = 0; k < 256; k++)
		if (tree_fits(oz, k >> 4, k & 15, cols[k], &base)) tree_trunk(chunk, k >> 4, k & 15, base, cols[k]);
}

// The whole region on the host (mm_gen_terrain_host of both libraries).  false: bad arguments.
static inline bool gen_region_host(uint64_t seed, long long cx0, long long cy0, long long cz0, uint32_t nx, uint32_t ny,
                                   uint32_t nz, uint8_t *chunks, long long *offs) {
	if (!chunks || !offs || !nx || !ny || !nz) return false;
	const Region rg = {seed, cx0, cy0, cz0, nx, ny, nz};
	const uint64_t n = (uint64_t)nx * ny * nz;
	for (uint64_t i = 0; i < n; i++) gen_chunk_host(rg, (uint32_t)i, chunks + i * 4096, offs + 3 * i);
	return true;
}

}  // namespace tg
