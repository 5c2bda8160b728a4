// terrain_host.cpp -- libmimas_terrain_host.so: the synthetic mapgen look-alike (terrain_gen.h) for hosts
// without CUDA.  Input generation only: bench.py's CPU arms and the CPU tests build their chunks with it, so
// that they never map the CUDA product library.  Same integer code, same bytes as mm_gen_terrain_device.
#include <stdint.h>

#include "terrain_gen.h"

extern "C" int mm_gen_terrain_host(uint64_t seed, int64_t cx0, int64_t cy0, int64_t cz0, uint32_t nx, uint32_t ny,
                                   uint32_t nz, uint8_t *chunks, int64_t *offs) {
	return tg::gen_region_host(seed, cx0, cy0, cz0, nx, ny, nz, chunks, (long long *)offs) ? 0 : 1;
}
