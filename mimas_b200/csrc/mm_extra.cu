// mm_extra.cu -- the remaining entry points: push_block cube emitter and the synthetic
// terrain generator (host and device builds of the same integer code).
#include <cuda_runtime.h>
#include <string.h>

#include <vector>

#include "mm_internal.h"
#include "terrain_gen.h"

namespace {

#define fail mm_fail

// ---- push_block (mimas-meshgen/lib.rs:183-219) -------------------------------------------
// One thread per (cube, face).  Face f of cube i lands at quad slot
// base[i] + popc(~blocked & ((1 << f) - 1)), i.e. faces keep their source order
// (lib.rs:196-218) and blocked faces are skipped.
__global__ void push_blocks_kernel(const mm_cube *cubes, const uint32_t *base, uint32_t n, unsigned char *out) {
	const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t i = t / 6u, f = t % 6u;
	if (i >= n) return;
	const mm_cube c = cubes[i];
	const uint32_t open = ~(uint32_t)c.blocked_mask & 63u;
	if (!((open >> f) & 1u)) return;
	const uint32_t slot = base[i] + (uint32_t)__popc(open & ((1u << f) - 1u));
	const float x = c.pos[0], y = c.pos[1], z = c.pos[2], siz = c.siz;
	mm::QuadArgs a;
	a.x = x; a.y = y; a.z = z;
	switch (f) {
	case 0: a.xsd = siz; a.ysd = 0.f; a.yd = siz; a.zd = 0.f; a.tex = c.id_bottom; a.rev = false; break;               // lib.rs:197
	case 1: a.xsd = siz; a.ysd = 0.f; a.yd = 0.f; a.zd = siz; a.tex = c.id_sides; a.rev = true; break;                 // lib.rs:201
	case 2: a.xsd = 0.f; a.ysd = siz; a.yd = 0.f; a.zd = siz; a.tex = c.id_sides; a.rev = false; break;                // lib.rs:205
	case 3: a.z = mm::fadd(z, siz); a.xsd = siz; a.ysd = 0.f; a.yd = siz; a.zd = 0.f; a.tex = c.id_top; a.rev = true; break;   // lib.rs:209
	case 4: a.y = mm::fadd(y, siz); a.xsd = siz; a.ysd = 0.f; a.yd = 0.f; a.zd = siz; a.tex = c.id_sides; a.rev = false; break; // lib.rs:213
	default: a.x = mm::fadd(x, siz); a.xsd = 0.f; a.ysd = siz; a.yd = 0.f; a.zd = siz; a.tex = c.id_sides; a.rev = true; break; // lib.rs:217
	}
	a.us = mm::fadd(a.xsd, a.ysd);
	a.vt = mm::fadd(a.yd, a.zd);
	mm::write_quad(out + (size_t)slot * mm::kQuadBytes, a);
}

// ---- terrain generator ---------------------------------------------------------------------
using tg::Region;
using tg::tree_leaves;
using tg::tree_trunk;

// one CTA of 256 threads per chunk, one thread per column
__global__ void __launch_bounds__(256) gen_terrain_kernel(Region rg, uint32_t n, uint8_t *chunks, long long *offs) {
	__shared__ __align__(16) uint8_t sc[4096];
	for (uint32_t i = blockIdx.x; i < n; i += gridDim.x) {
		const long long cx = rg.cx0 + (long long)(i % rg.nx), cy = rg.cy0 + (long long)((i / rg.nx) % rg.ny),
		                cz = rg.cz0 + (long long)(i / (rg.nx * rg.ny));
		const long long ox = cx * 16, oy = cy * 16, oz = cz * 16;
		const int k = threadIdx.x, x = k >> 4, y = k & 15;
		if (k < 3) offs[3 * (size_t)i + k] = k == 0 ? ox : (k == 1 ? oy : oz);
		const tg::Column ci = tg::column_info(rg.seed, ox + x, oy + y);
		uint8_t col[16];
		tg::gen_column(rg.seed, ox, oy, oz, x, y, ci, col);
		for (int z = 0; z < 16; z++) sc[k * 16 + z] = col[z];
		__syncthreads();
		int base = 0;
		const bool tree = tg::tree_fits(oz, x, y, ci, &base);
		if (tree) tree_leaves(sc, x, y, base, ci);
		__syncthreads();
		if (tree) tree_trunk(sc, x, y, base, ci);
		__syncthreads();
		reinterpret_cast<uint4 *>(chunks + (size_t)i * 4096)[k] = reinterpret_cast<const uint4 *>(sc)[k];
		__syncthreads();
	}
}

// ---- halo slabs of a resident box of chunks (extension, see mm_mesh_batch) ------------------------
// One CTA of 96 threads per chunk: thread t writes row t & 15 (16 ids) of slab t >> 4.  The x and y slabs
// are 16-byte rows of the neighbour chunk; the z slabs gather its z = 15 / z = 0 layer (stride 16; the
// box is L2-resident when it was just generated or uploaded).
__global__ void __launch_bounds__(96) gather_halos_kernel(const uint8_t *chunks, uint32_t nx, uint32_t ny, uint32_t nz,
                                                          uint32_t n, uint32_t outside, uint8_t *halo) {
	for (uint32_t i = blockIdx.x; i < n; i += gridDim.x) {
		const uint32_t cx = i % nx, cy = (i / nx) % ny, cz = i / (nx * ny);
		const int t = threadIdx.x, slab = t >> 4, r = t & 15;
		const int axis = slab >> 1;
		const bool up = slab & 1;
		const uint32_t pos = axis == 0 ? cx : axis == 1 ? cy : cz, lim = axis == 0 ? nx : axis == 1 ? ny : nz;
		uint4 v = make_uint4(outside, outside, outside, outside);
		if (up ? pos + 1 < lim : pos > 0) {
			const size_t stride = axis == 0 ? 1 : axis == 1 ? nx : (size_t)nx * ny;
			const uint8_t *nb = chunks + (up ? i + stride : i - stride) * 4096u;
			const int edge = up ? 0 : 15;  // the neighbour's layer that touches this chunk
			if (axis == 0) v = __ldg(reinterpret_cast<const uint4 *>(nb + edge * 256 + r * 16));       // (edge, y = r, z)
			else if (axis == 1) v = __ldg(reinterpret_cast<const uint4 *>(nb + r * 256 + edge * 16));  // (x = r, edge, z)
			else {                                                                                       // (x = r, y, edge)
				uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
				for (int y = 0; y < 16; y++) w[y >> 2] |= (uint32_t)__ldg(nb + r * 256 + y * 16 + edge) << (8 * (y & 3));
				v = make_uint4(w[0], w[1], w[2], w[3]);
			}
		}
		reinterpret_cast<uint4 *>(halo + (size_t)i * MM_HALO_BYTES)[t] = v;
	}
}

}  // namespace

extern "C" {

int mm_push_blocks(mm_ctx *c, const mm_cube *cubes, uint32_t n, mm_vertex *out, uint64_t *n_out) {
	if (!c || !n_out || (n && (!cubes || !out))) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	*n_out = 0;
	if (n == 0) return MM_OK;
	// per-cube base slots: a host-side running sum (UI meshes are <= a few cubes)
	std::vector<uint32_t> base(n);
	uint32_t total = 0;
	for (uint32_t i = 0; i < n; i++) {
		base[i] = total;
		total += (uint32_t)__builtin_popcount(~(uint32_t)cubes[i].blocked_mask & 63u);
	}
	mm_cube *d_c = nullptr;
	uint32_t *d_b = nullptr;
	unsigned char *d_o = nullptr;
	int rc = MM_OK;
	auto body = [&]() -> int {
		MM_CUDA(c, cudaMalloc(&d_c, sizeof(mm_cube) * (size_t)n));
		MM_CUDA(c, cudaMalloc(&d_b, sizeof(uint32_t) * (size_t)n));
		MM_CUDA(c, cudaMalloc(&d_o, (size_t)total * mm::kQuadBytes + 16));
		MM_CUDA(c, cudaMemcpyAsync(d_c, cubes, sizeof(mm_cube) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
		MM_CUDA(c, cudaMemcpyAsync(d_b, base.data(), sizeof(uint32_t) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
		const uint32_t threads = n * 6u;
		push_blocks_kernel<<<(threads + 127u) / 128u, 128, 0, c->stream>>>(d_c, d_b, n, d_o);
		MM_CUDA(c, cudaGetLastError());
		if (total) MM_CUDA(c, cudaMemcpyAsync(out, d_o, (size_t)total * mm::kQuadBytes, cudaMemcpyDeviceToHost, c->stream));
		MM_CUDA(c, cudaStreamSynchronize(c->stream));
		return MM_OK;
	};
	rc = body();
	cudaFree(d_c); cudaFree(d_b); cudaFree(d_o);
	if (rc == MM_OK) *n_out = (uint64_t)total * 6u;
	return rc;
}

int mm_gen_terrain_host(uint64_t seed, int64_t cx0, int64_t cy0, int64_t cz0, uint32_t nx, uint32_t ny, uint32_t nz,
                        uint8_t *chunks, int64_t *offs) {
	if (!chunks || !offs || !nx || !ny || !nz) return MM_ERR_INVALID;
	return tg::gen_region_host(seed, cx0, cy0, cz0, nx, ny, nz, chunks, (long long *)offs) ? MM_OK : MM_ERR_INVALID;
}

int mm_gen_terrain_device(mm_ctx *c, uint64_t seed, int64_t cx0, int64_t cy0, int64_t cz0, uint32_t nx, uint32_t ny,
                          uint32_t nz, uint8_t *d_chunks, int64_t *d_offs) {
	if (!c || !d_chunks || !d_offs || !nx || !ny || !nz) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	const uint64_t n64 = (uint64_t)nx * ny * nz;
	if (n64 > 0x7fffffffull) return fail(c, MM_ERR_INVALID, "region too large");
	const Region rg = {seed, cx0, cy0, cz0, nx, ny, nz};
	const uint32_t n = (uint32_t)n64;
	const uint32_t grid = n < (uint32_t)c->sm_count * 8u ? n : (uint32_t)c->sm_count * 8u;
	gen_terrain_kernel<<<grid, 256, 0, c->stream>>>(rg, n, d_chunks, (long long *)d_offs);
	MM_CUDA(c, cudaGetLastError());
	return MM_OK;
}

int mm_gather_halos(mm_ctx *c, const uint8_t *d_chunks, uint32_t nx, uint32_t ny, uint32_t nz, uint8_t outside_id,
                    uint8_t *d_halo) {
	if (!c || !d_chunks || !d_halo || !nx || !ny || !nz) return MM_ERR_INVALID;
	if (((uintptr_t)d_halo | (uintptr_t)d_chunks) & 15u) return fail(c, MM_ERR_INVALID, "chunks and halo must be 16-byte aligned");
	MM_CUDA(c, cudaSetDevice(c->device));
	const uint64_t n64 = (uint64_t)nx * ny * nz;
	if (n64 > 0x7fffffffull) return fail(c, MM_ERR_INVALID, "region too large");
	const uint32_t n = (uint32_t)n64;
	const uint32_t grid = n < (uint32_t)c->sm_count * 16u ? n : (uint32_t)c->sm_count * 16u;
	const uint32_t o = outside_id * 0x01010101u;
	gather_halos_kernel<<<grid, 96, 0, c->stream>>>(d_chunks, nx, ny, nz, n, o, d_halo);
	MM_CUDA(c, cudaGetLastError());
	return MM_OK;
}

}  // extern "C"
