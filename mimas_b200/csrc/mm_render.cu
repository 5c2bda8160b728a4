// mm_render.cu -- headless consumer of the vertex stream: chunk culling
// (mimas-client/client.rs:679-696 with the sphere/frustum test of frustum_query 0.1.2) and the vertex
// shader (mimas-client/vertex-shader.glsl:18-27) as CUDA kernels.  GL is unavailable on this box; the
// shaded vertex buffer in device memory is the headless stand-in for what the rasteriser would be fed.
// Float operations are ordered exactly as oracle/render_oracle.c orders them (separately rounded
// products and sums, no FMA: the library is built with -fmad=false), so the two agree bit for bit.
#include <cuda_runtime.h>
#include <string.h>

#include <algorithm>

#include "mm_internal.h"

namespace {

#define fail mm_fail
constexpr int kCullThreads = 1024;
constexpr int kShadeThreads = 256;

struct Planes {
	float p[6][4];
};

// frustum_query: clip = modelview * projection on column-major arrays; planes right, left, bottom, top, far, near
void make_planes(const mm_camera &cam, Planes &pl) {
	const float *modl = cam.vmatrix, *proj = cam.pmatrix;
	volatile float clip[16];  // volatile: every product and sum rounded to f32, as the oracle does with -ffp-contract=off
	for (int r = 0; r < 4; r++)
		for (int c = 0; c < 4; c++) {
			volatile float a = modl[4 * r + 0] * proj[0 + c], b = modl[4 * r + 1] * proj[4 + c];
			volatile float d = modl[4 * r + 2] * proj[8 + c], e = modl[4 * r + 3] * proj[12 + c];
			volatile float s = a + b;
			s = s + d;
			s = s + e;
			clip[4 * r + c] = s;
		}
	static const int comp[6] = {0, 0, 1, 1, 2, 2};
	static const float sign[6] = {-1.0f, 1.0f, 1.0f, -1.0f, -1.0f, 1.0f};
	for (int p = 0; p < 6; p++) {
		volatile float v[4];
		for (int k = 0; k < 4; k++) {
			volatile float m = sign[p] * clip[4 * k + comp[p]];
			v[k] = clip[4 * k + 3] + m;
		}
		volatile float x = v[0] * v[0], y = v[1] * v[1], z = v[2] * v[2];
		volatile float s = x + y;
		s = s + z;
		const float t = sqrtf(s);
		for (int k = 0; k < 4; k++) pl.p[p][k] = v[k] / t;
	}
}

struct CullParams {
	const long long *offs;
	uint32_t n;
	Planes pl;
	float pos[3], viewing_range, radius;
	const mm::Range *ranges_i, *ranges_t;  // may be NULL (mm_cull_chunks): vertex counts are then zero
	unsigned char *flag;                   // n
	uint32_t *bs_cnt;                      // per block of kCullThreads chunks: visible chunks
	unsigned long long *bs_vi, *bs_vt;     // .. and their vertices; after scan_blocks: exclusive prefixes
	uint32_t *draw;                        // n
	unsigned long long *first_i, *first_t; // per drawn chunk: where its shaded vertices start
	unsigned long long *totals;            // [0] drawn chunks, [1] vertices i, [2] vertices t (mapped host memory)
	unsigned long long *totals_dev;        // the same three words in device memory, for the draw kernel
};

__device__ __forceinline__ bool chunk_visible(const CullParams &p, uint32_t i) {
	const long long ox = p.offs[3 * (size_t)i], oy = p.offs[3 * (size_t)i + 1], oz = p.offs[3 * (size_t)i + 2];
	// viewing range on the chunk origin, client.rs:681-685
	const float dx = __fsub_rn(__ll2float_rn(ox), p.pos[0]), dy = __fsub_rn(__ll2float_rn(oy), p.pos[1]),
	            dz = __fsub_rn(__ll2float_rn(oz), p.pos[2]);
	const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
	if (__fsqrt_rn(d2) > p.viewing_range) return false;
	// sphere around the chunk centre against the six planes, client.rs:686-693
	const float cx = __ll2float_rn(ox + 8), cy = __ll2float_rn(oy + 8), cz = __ll2float_rn(oz + 8);
#pragma unroll
	for (int k = 0; k < 6; k++) {
		const float d = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.pl.p[k][0], cx), __fmul_rn(p.pl.p[k][1], cy)),
		                                    __fmul_rn(p.pl.p[k][2], cz)), p.pl.p[k][3]);
		if (d <= -p.radius) return false;
	}
	return true;
}

// block-wide exclusive scan of one u32 per thread (kCullThreads threads); returns the exclusive prefix, *total = block sum
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *warp_sums, uint32_t *total) {
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
		if (lane >= d) inc += o;
	}
	__syncthreads();  // warp_sums may still be read from the previous scan
	if (lane == 31) warp_sums[warp] = inc;
	__syncthreads();
	uint32_t w = warp_sums[lane];  // kCullThreads / 32 == 32 warps
	uint32_t winc = w;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t o = __shfl_up_sync(0xffffffffu, winc, d);
		if (lane >= d) winc += o;
	}
	*total = __shfl_sync(0xffffffffu, winc, 31);
	return __shfl_sync(0xffffffffu, winc - w, warp) + inc - v;
}

__global__ void __launch_bounds__(kCullThreads) cull_flags_kernel(const CullParams p) {
	__shared__ uint32_t ws[32];
	const uint32_t i = blockIdx.x * kCullThreads + threadIdx.x;
	const bool vis = i < p.n && chunk_visible(p, i);
	if (i < p.n) p.flag[i] = vis ? 1 : 0;
	const uint32_t vi = vis && p.ranges_i ? p.ranges_i[i].vertex_count : 0u, vt = vis && p.ranges_t ? p.ranges_t[i].vertex_count : 0u;
	uint32_t tc, ti, tt;
	block_excl_scan(vis ? 1u : 0u, ws, &tc);
	block_excl_scan(vi, ws, &ti);  // <= 1024 * 98304 vertices: fits 32 bits
	block_excl_scan(vt, ws, &tt);
	if (threadIdx.x == 0) {
		p.bs_cnt[blockIdx.x] = tc;
		p.bs_vi[blockIdx.x] = ti;
		p.bs_vt[blockIdx.x] = tt;
	}
}

// one block: exclusive prefixes of the per-block sums, totals to the host
__global__ void __launch_bounds__(kCullThreads) scan_blocks_kernel(const CullParams p, uint32_t nb) {
	__shared__ unsigned long long carry[3];
	__shared__ unsigned long long wsum[3][32];
	if (threadIdx.x < 3) carry[threadIdx.x] = 0;
	__syncthreads();
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (uint32_t b0 = 0; b0 < nb; b0 += kCullThreads) {
		const uint32_t b = b0 + threadIdx.x;
		unsigned long long v[3] = {b < nb ? p.bs_cnt[b] : 0ull, b < nb ? p.bs_vi[b] : 0ull, b < nb ? p.bs_vt[b] : 0ull};
		unsigned long long inc[3];
#pragma unroll
		for (int k = 0; k < 3; k++) {
			inc[k] = v[k];
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const unsigned long long o = __shfl_up_sync(0xffffffffu, inc[k], d);
				if (lane >= d) inc[k] += o;
			}
			if (lane == 31) wsum[k][warp] = inc[k];
		}
		__syncthreads();
		unsigned long long base[3];
#pragma unroll
		for (int k = 0; k < 3; k++) {
			unsigned long long s = carry[k];
			for (int w = 0; w < warp; w++) s += wsum[k][w];
			base[k] = s + inc[k] - v[k];
		}
		if (b < nb) {
			p.bs_cnt[b] = (uint32_t)base[0];
			p.bs_vi[b] = base[1];
			p.bs_vt[b] = base[2];
		}
		__syncthreads();
		if (threadIdx.x == kCullThreads - 1) {
#pragma unroll
			for (int k = 0; k < 3; k++) carry[k] = base[k] + v[k];
		}
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		p.totals_dev[0] = carry[0];
		p.totals_dev[1] = carry[1];
		p.totals_dev[2] = carry[2];
		p.totals[0] = carry[0];
		p.totals[1] = carry[1];
		p.totals[2] = carry[2];
		__threadfence_system();
	}
}

__global__ void __launch_bounds__(kCullThreads) cull_write_kernel(const CullParams p) {
	__shared__ uint32_t ws[32];
	const uint32_t i = blockIdx.x * kCullThreads + threadIdx.x;
	const bool vis = i < p.n && p.flag[i];
	const uint32_t vi = vis && p.ranges_i ? p.ranges_i[i].vertex_count : 0u, vt = vis && p.ranges_t ? p.ranges_t[i].vertex_count : 0u;
	uint32_t t;
	const uint32_t rank = block_excl_scan(vis ? 1u : 0u, ws, &t);
	const uint32_t ei = block_excl_scan(vi, ws, &t), et = block_excl_scan(vt, ws, &t);
	if (vis) {
		const uint32_t k = p.bs_cnt[blockIdx.x] + rank;
		p.draw[k] = i;
		p.first_i[k] = p.bs_vi[blockIdx.x] + ei;
		p.first_t[k] = p.bs_vt[blockIdx.x] + et;
	}
}

// ---- vertex shader ---------------------------------------------------------------------------------
struct Mats {
	float v[16], p[16];
};

__device__ __forceinline__ void mat_vec(const float *m, float x, float y, float z, float w, float *out) {
#pragma unroll
	for (int r = 0; r < 4; r++)
		out[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[r], x), __fmul_rn(m[4 + r], y)), __fmul_rn(m[8 + r], z)),
		                   __fmul_rn(m[12 + r], w));
}

// vertex-shader.glsl:18-27 for one vertex
__device__ __forceinline__ void shade_vertex(const Mats &mt, const mm_vertex *in, mm_shaded_vertex *out) {
	const uint32_t *w = reinterpret_cast<const uint32_t *>(in);
	const uint32_t w0 = __ldg(w);
	const float u = __uint_as_float(__ldg(w + 1)), v = __uint_as_float(__ldg(w + 2));
	const float px = __uint_as_float(__ldg(w + 3)), py = __uint_as_float(__ldg(w + 4)), pz = __uint_as_float(__ldg(w + 5));
	const float nx0 = __uint_as_float(__ldg(w + 6)), ny0 = __uint_as_float(__ldg(w + 7)), nz0 = __uint_as_float(__ldg(w + 8));
	float vp[4], cl[4];
	mat_vec(mt.v, px, py, pz, 1.0f, vp);          // vposition = vmatrix * vec4(position, 1.0)
	mat_vec(mt.p, vp[0], vp[1], vp[2], vp[3], cl);  // gl_Position = pmatrix * vposition
	const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx0, nx0), __fmul_rn(ny0, ny0)), __fmul_rn(nz0, nz0)));
	const float nx = __fdiv_rn(nx0, len), ny = __fdiv_rn(ny0, len), nz = __fdiv_rn(nz0, len);
	const float l = __fdiv_rn(1.0f, __fsqrt_rn(2.0f));  // normalize(vec3(0, -1, -1)) = (0, -l, -l)
	const float da = __fadd_rn(__fadd_rn(__fmul_rn(0.0f, nx), __fmul_rn(-l, ny)), __fmul_rn(-l, nz));
	const float db = __fadd_rn(__fadd_rn(__fmul_rn(0.0f, nx), __fmul_rn(l, ny)), __fmul_rn(l, nz));
	const float vlamb = __fadd_rn(fmaxf(da, 0.2f), fmaxf(db, 0.2f));
	float4 *o = reinterpret_cast<float4 *>(out);
	o[0] = make_float4(cl[0], cl[1], cl[2], cl[3]);
	o[1] = make_float4(vp[0], vp[1], vp[2], vp[3]);
	o[2] = make_float4(vlamb, u, v, __uint_as_float(w0 & 0xffffu));
}

__global__ void __launch_bounds__(kShadeThreads) vertex_stage_kernel(const mm_vertex *in, unsigned long long n, const Mats mt,
                                                                     mm_shaded_vertex *out) {
	for (unsigned long long i = (unsigned long long)blockIdx.x * kShadeThreads + threadIdx.x; i < n;
	     i += (unsigned long long)gridDim.x * kShadeThreads)
		shade_vertex(mt, in + i, out + i);
}

// one block per drawn chunk (persistent grid): the chunk's intransparent vertices go to first_i[k], its
// transparent ones behind all intransparent vertices of the frame (client.rs:699-701)
struct DrawParams {
	const uint32_t *draw;
	const unsigned long long *first_i, *first_t, *totals;
	const mm::Range *ranges_i, *ranges_t;
	const mm_vertex *vi, *vt;
	mm_shaded_vertex *out;
	unsigned long long cap;  // vertices `out` can hold
};
__global__ void __launch_bounds__(kShadeThreads) draw_kernel(const DrawParams p, const Mats mt) {
	const uint32_t nd = (uint32_t)p.totals[0];
	const unsigned long long tot_i = p.totals[1], tot_t = p.totals[2];
	if (tot_i + tot_t > p.cap) return;  // the host grows the buffer and runs the frame again
	for (uint32_t k = blockIdx.x; k < nd; k += gridDim.x) {
		const uint32_t c = p.draw[k];
		const mm::Range ri = p.ranges_i[c], rt = p.ranges_t[c];
		const mm_vertex *si = p.vi + ri.first_vertex;
		mm_shaded_vertex *di = p.out + p.first_i[k];
		for (uint32_t v = threadIdx.x; v < ri.vertex_count; v += kShadeThreads) shade_vertex(mt, si + v, di + v);
		const mm_vertex *st = p.vt + rt.first_vertex;
		mm_shaded_vertex *dt = p.out + tot_i + p.first_t[k];
		for (uint32_t v = threadIdx.x; v < rt.vertex_count; v += kShadeThreads) shade_vertex(mt, st + v, dt + v);
	}
}

bool is_device_ptr(const void *p) {
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// scratch layout for n chunks; fills the pointer fields of p
int setup_cull(mm_ctx *c, uint32_t n, CullParams &p) {
	const uint32_t nb = (n + kCullThreads - 1) / kCullThreads;
	const size_t o_flag = 0, o_draw = align256(o_flag + n), o_fi = align256(o_draw + 4 * (size_t)n), o_ft = align256(o_fi + 8 * (size_t)n),
	             o_bc = align256(o_ft + 8 * (size_t)n), o_bi = align256(o_bc + 4 * (size_t)nb), o_bt = align256(o_bi + 8 * (size_t)nb),
	             total = align256(o_bt + 8 * (size_t)nb) + 256;
	if (total > c->cull_bytes) {
		MM_CUDA(c, cudaStreamSynchronize(c->stream));
		if (c->s_draw) MM_CUDA(c, cudaStreamSynchronize(c->s_draw));
		cudaFree(c->d_cull);
		c->d_cull = nullptr;
		c->cull_bytes = 0;
		MM_CUDA(c, cudaMalloc(&c->d_cull, total));
		c->cull_bytes = total;
	}
	p.flag = c->d_cull + o_flag;
	p.draw = (uint32_t *)(c->d_cull + o_draw);
	p.first_i = (unsigned long long *)(c->d_cull + o_fi);
	p.first_t = (unsigned long long *)(c->d_cull + o_ft);
	p.bs_cnt = (uint32_t *)(c->d_cull + o_bc);
	p.bs_vi = (unsigned long long *)(c->d_cull + o_bi);
	p.bs_vt = (unsigned long long *)(c->d_cull + o_bt);
	p.totals_dev = (unsigned long long *)(c->d_cull + total - 256);  // the spare 256 bytes at the end
	return MM_OK;
}

int enqueue_cull(mm_ctx *c, const CullParams &p, cudaStream_t stream = nullptr) {
	if (!stream) stream = c->stream;
	const uint32_t nb = (p.n + kCullThreads - 1) / kCullThreads;
	cull_flags_kernel<<<nb, kCullThreads, 0, stream>>>(p);
	scan_blocks_kernel<<<1, kCullThreads, 0, stream>>>(p, nb);
	cull_write_kernel<<<nb, kCullThreads, 0, stream>>>(p);
	MM_CUDA(c, cudaGetLastError());
	return MM_OK;
}

void fill_camera(const mm_camera &cam, CullParams &p, Mats &mt) {
	make_planes(cam, p.pl);
	memcpy(p.pos, cam.pos, sizeof(p.pos));
	p.viewing_range = cam.viewing_range;
	p.radius = 16.0f * sqrtf(3.0f);  // CHUNKSIZE as f32 * 3.0_f32.sqrt(), client.rs:690
	memcpy(mt.v, cam.vmatrix, sizeof(mt.v));
	memcpy(mt.p, cam.pmatrix, sizeof(mt.p));
}

}  // namespace

namespace {
int ensure_draw_totals(mm_ctx *c) {
	if (c->h_draw_totals) return MM_OK;
	MM_CUDA(c, cudaHostAlloc(&c->h_draw_totals, 4 * sizeof(unsigned long long), cudaHostAllocMapped));
	MM_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_h_draw_totals, c->h_draw_totals, 0));
	if (!c->ev2) {  // mm_raster_frame may have made them already
		MM_CUDA(c, cudaEventCreate(&c->ev2));
		MM_CUDA(c, cudaEventCreate(&c->ev3));
	}
	return MM_OK;
}
}  // namespace

extern "C" {

int mm_vertex_stage(mm_ctx *c, const mm_vertex *d_in, uint64_t n, const mm_camera *cam, mm_shaded_vertex *d_out) {
	if (!c || !cam || (n && (!d_in || !d_out))) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	if (n == 0) return MM_OK;
	Mats mt;
	memcpy(mt.v, cam->vmatrix, sizeof(mt.v));
	memcpy(mt.p, cam->pmatrix, sizeof(mt.p));
	const unsigned long long blocks = (n + kShadeThreads - 1) / kShadeThreads;
	const unsigned grid = (unsigned)std::min<unsigned long long>(blocks, (unsigned long long)c->sm_count * 16);
	vertex_stage_kernel<<<grid, kShadeThreads, 0, c->stream>>>(d_in, n, mt, d_out);
	MM_CUDA(c, cudaGetLastError());
	return MM_OK;
}

int mm_cull_chunks(mm_ctx *c, const int64_t *offs, uint32_t n, const mm_camera *cam, uint32_t *h_draw, uint32_t *n_drawn) {
	if (!c || !cam || !n_drawn || (n && (!offs || !h_draw))) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	*n_drawn = 0;
	if (n == 0) return MM_OK;
	int rc;
	if ((rc = ensure_draw_totals(c))) return rc;
	CullParams p;
	memset(&p, 0, sizeof(p));
	Mats mt;
	fill_camera(*cam, p, mt);
	if ((rc = setup_cull(c, n, p))) return rc;
	long long *d_offs = nullptr;
	if (is_device_ptr(offs)) {
		p.offs = (const long long *)offs;
	} else {
		MM_CUDA(c, cudaMalloc(&d_offs, (size_t)n * 24));
		MM_CUDA(c, cudaMemcpyAsync(d_offs, offs, (size_t)n * 24, cudaMemcpyHostToDevice, c->stream));
		p.offs = d_offs;
	}
	p.n = n;
	p.totals = c->d_h_draw_totals;
	rc = enqueue_cull(c, p);
	if (rc == MM_OK) {
		cudaError_t e = cudaStreamSynchronize(c->stream);
		if (e == cudaSuccess) {
			*n_drawn = (uint32_t)c->h_draw_totals[0];
			if (*n_drawn) e = cudaMemcpy(h_draw, p.draw, sizeof(uint32_t) * (size_t)*n_drawn, cudaMemcpyDeviceToHost);
		}
		if (e != cudaSuccess) rc = fail(c, MM_ERR_CUDA, "mm_cull_chunks: %s", cudaGetErrorString(e));
	}
	if (d_offs) cudaFree(d_offs);
	return rc;
}

// One frame over the batch a ring slot holds, on `stream`: enqueue (cull + vertex shader), then collect (wait, grow the
// shaded buffer and run again if it was too small, read the totals).  One frame in flight per context.
static int draw_enqueue(mm_ctx *c, const mm_slot *sl, cudaStream_t stream, const mm_camera *cam) {
	c->draw_slot = nullptr;
	if (sl->last_n == 0) {
		c->draw_slot = sl;
		c->draw_stream = stream;
		return MM_OK;
	}
	int rc;
	if ((rc = ensure_draw_totals(c))) return rc;
	CullParams p;
	memset(&p, 0, sizeof(p));
	Mats mt;
	fill_camera(*cam, p, mt);
	if ((rc = setup_cull(c, sl->last_n, p))) return rc;
	p.offs = sl->last_offs;
	p.n = sl->last_n;
	p.ranges_i = sl->d_ranges_i;
	p.ranges_t = sl->d_ranges_t;
	p.totals = c->d_h_draw_totals;
	DrawParams d;
	d.draw = p.draw; d.first_i = p.first_i; d.first_t = p.first_t; d.totals = p.totals_dev;  // device copy: no PCIe read per CTA
	d.ranges_i = sl->d_ranges_i; d.ranges_t = sl->d_ranges_t;
	d.vi = (const mm_vertex *)sl->d_out_i; d.vt = (const mm_vertex *)sl->d_out_t;
	d.out = (mm_shaded_vertex *)c->d_shaded;
	d.cap = c->shaded_bytes / sizeof(mm_shaded_vertex);
	MM_CUDA(c, cudaEventRecord(c->ev2, stream));
	if ((rc = enqueue_cull(c, p, stream))) return rc;
	draw_kernel<<<c->sm_count * 8, kShadeThreads, 0, stream>>>(d, mt);
	MM_CUDA(c, cudaGetLastError());
	MM_CUDA(c, cudaEventRecord(c->ev3, stream));
	c->draw_slot = sl;
	c->draw_stream = stream;
	c->draw_cam = *cam;
	c->draw_list = p.draw;
	return MM_OK;
}

static int draw_collect(mm_ctx *c, mm_draw_out *out) {
	const mm_slot *sl = c->draw_slot;
	if (!sl) return fail(c, MM_ERR_NO_BATCH, "no frame in flight");
	memset(out, 0, sizeof(*out));
	out->n_chunks = sl->last_n;
	c->draw_slot = nullptr;
	if (sl->last_n == 0) return MM_OK;
	for (int attempt = 0;; attempt++) {
		MM_CUDA(c, cudaStreamSynchronize(c->draw_stream));
		const unsigned long long need = c->h_draw_totals[1] + c->h_draw_totals[2];
		if (need <= c->shaded_bytes / sizeof(mm_shaded_vertex)) break;
		if (attempt >= 1) return fail(c, MM_ERR_CUDA, "shaded vertex buffer still too small after growth");
		cudaFree(c->d_shaded);
		c->d_shaded = nullptr;
		c->shaded_bytes = 0;
		const size_t bytes = align256((size_t)(need + need / 8) * sizeof(mm_shaded_vertex));
		MM_CUDA(c, cudaMalloc(&c->d_shaded, bytes));
		c->shaded_bytes = bytes;
		const mm_camera cam = c->draw_cam;
		int rc;
		if ((rc = draw_enqueue(c, sl, c->draw_stream, &cam))) return rc;
		c->draw_slot = nullptr;
	}
	MM_CUDA(c, cudaEventElapsedTime(&out->kernel_ms, c->ev2, c->ev3));
	out->d_shaded = (const mm_shaded_vertex *)c->d_shaded;
	out->d_draw_list = c->draw_list;
	out->n_drawn = (uint32_t)c->h_draw_totals[0];
	out->n_shaded_i = c->h_draw_totals[1];
	out->n_shaded_t = c->h_draw_totals[2];
	return MM_OK;
}

int mm_draw_batch(mm_ctx *c, const mm_camera *cam, mm_draw_out *out) {
	if (!c || !cam || !out) return MM_ERR_INVALID;
	if (!c->cur->finished) return fail(c, MM_ERR_NO_BATCH, "no finished device batch to draw");
	MM_CUDA(c, cudaSetDevice(c->device));
	int rc;
	if ((rc = draw_enqueue(c, c->cur, c->stream, cam))) return rc;
	return draw_collect(c, out);
}

int mm_draw_ticket_async(mm_ctx *c, uint64_t ticket, const mm_camera *cam) {
	if (!c || !cam || !ticket) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	if (c->draw_slot) return fail(c, MM_ERR_INVALID, "a frame is in flight: collect it first (mm_draw_collect)");
	int rc;
	if ((rc = mm_batch_wait(c, ticket, nullptr))) return rc;  // polls the batch's completion word: no stream synchronisation
	return draw_enqueue(c, &c->ring[ticket % c->ring_slots], c->s_draw, cam);
}

int mm_draw_collect(mm_ctx *c, mm_draw_out *out) {
	if (!c || !out) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	return draw_collect(c, out);
}

int mm_draw_ticket(mm_ctx *c, uint64_t ticket, const mm_camera *cam, mm_draw_out *out) {
	if (!out) return MM_ERR_INVALID;
	const int rc = mm_draw_ticket_async(c, ticket, cam);
	return rc ? rc : mm_draw_collect(c, out);
}

}  // extern "C"
