// mm_raster.cu -- headless stand-in for what GL does with the shaded vertices (the
// optional part): primitive assembly, clipping, back-face culling, rasterisation, depth test and the
// fragment shader (mimas-client/fragment-shader.glsl:14-24) under the draw state of
// mimas-client/client.rs:661-675, into an RGBA8 image in device memory.  GL is unavailable on this box.
//
// Parity: unpinned against the reference (GL leaves rasterisation rules, depth quantisation and float order
// to the driver; the reference has no images).  The kernels are bit-exact to oracle/raster_oracle.c, which
// fixes one choice of those rules and lists the deliberate simplifications (nearest texel at level 0 for
// minification too, no sRGB conversion, surviving fragments written unblended, depth ties to the smaller
// packed colour).  Every float operation below is written as that file writes it; the library is built
// with -fmad=false, IEEE division and square root.
//
// Shape: the frame buffer is one 64-bit key per pixel, depth24 << 32 | RGBA8; a fragment is one atomicMin
// (depth test IfLess + colour write in one step, independent of the order triangles arrive in).
//   raster_tris_kernel  one thread per triangle: clip, set up, and rasterise it if its bounding box is small;
//                       larger ones go to a list, cut into 64 x 64 pixel tiles (tiles wholly outside one
//                       edge are dropped), one queue entry per tile
//   raster_mid_kernel   one warp per listed triangle of at most 32 rows x 128 columns
//   raster_big_kernel   one CTA per queued (triangle, tile)
//                       both: thread <-> row solves the covered span of the row exactly from the three edge
//                       functions, a scan flattens the spans, thread <-> covered pixel shades
//   resolve_kernel      keys -> RGBA8 image (clear colour where nothing was drawn)
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>

#include "mm_internal.h"

namespace {

#define fail mm_fail
constexpr int kNAttr = 7;          // vlamb, u, v, vposition.xyzw
constexpr float kGuard = 16.0f;    // guard band: clip x, y at +-kGuard * w
constexpr int kMaxPoly = 12;
#ifndef MM_RASTER_SMALL
#define MM_RASTER_SMALL 32
#endif
constexpr int kSmallBox = MM_RASTER_SMALL;  // bounding boxes up to this many pixels are rasterised by the triangle's own thread
constexpr int kMidRows = 32, kMidCols = 128;  // up to this box a listed triangle is one warp's job
constexpr uint32_t kBigCap = 1u << 20;   // listed triangles
constexpr uint32_t kMidCap = 1u << 20;   // of which queued for the warp kernel
constexpr uint32_t kTileCap = 1u << 22;  // queued (triangle, tile) pairs
constexpr int kTileShift = 6;            // 64 x 64 pixels
constexpr uint32_t kNoTri = 0xffffffffu;
constexpr int kTriThreads = 128;
constexpr int kBigThreads = 128;

struct ClipV {
	float c[4];
	float a[kNAttr];
};
struct ScreenV {  // one corner of a set-up triangle
	int32_t X, Y;  // window coordinates, 1/256 pixel
	float z, invw;
	float a[kNAttr];
};
struct BigTri {
	ScreenV v[3];  // positive-area orientation
	uint32_t layer;
	uint32_t pad;
};

struct RasterParams {
	const mm_shaded_vertex *verts;
	uint64_t n_tris;
	uint32_t width, height;
	float fog_near, fog_far;
	const float *tex;  // [layers][ts][ts][4]
	uint32_t layers, ts;
	unsigned long long *fb;
	BigTri *big;
	uint2 *tiles;         // x = slot in big (kNoTri: empty entry), y = tile x | tile y << 16
	uint32_t *mid;        // slots in big
	uint32_t *big_count;  // [0] triangles listed, [1] tile entries queued, [2] mid entries queued
};

__device__ __forceinline__ float plane_dist(const ClipV &v, int p) {
	switch (p) {
	case 0: return v.c[2] + v.c[3];
	case 1: return v.c[3] - v.c[2];
	case 2: return v.c[0] + kGuard * v.c[3];
	case 3: return kGuard * v.c[3] - v.c[0];
	case 4: return v.c[1] + kGuard * v.c[3];
	default: return kGuard * v.c[3] - v.c[1];
	}
}

__device__ __forceinline__ ClipV cut(const ClipV &in, const ClipV &out, float di, float dq) {
	const float t = di / (di - dq);
	ClipV r;
#pragma unroll
	for (int k = 0; k < 4; k++) r.c[k] = in.c[k] + t * (out.c[k] - in.c[k]);
#pragma unroll
	for (int k = 0; k < kNAttr; k++) r.a[k] = in.a[k] + t * (out.a[k] - in.a[k]);
	return r;
}

__device__ int clip_poly(ClipV *poly, ClipV *tmp, int n) {
	for (int p = 0; p < 6 && n; p++) {
		int m = 0;
		for (int i = 0; i < n; i++) {
			const ClipV &a = poly[i], &b = poly[i + 1 == n ? 0 : i + 1];
			const float da = plane_dist(a, p), db = plane_dist(b, p);
			if (da >= 0.0f) {
				tmp[m++] = a;
				if (db < 0.0f) tmp[m++] = cut(a, b, da, db);
			} else if (db >= 0.0f) {
				tmp[m++] = cut(b, a, db, da);
			}
		}
		n = m;
		for (int i = 0; i < n; i++) poly[i] = tmp[i];
	}
	return n;
}

__device__ __forceinline__ ScreenV to_screen(const ClipV &v, uint32_t width, uint32_t height) {
	ScreenV s;
	s.invw = 1.0f / v.c[3];
	const float nx = v.c[0] * s.invw, ny = v.c[1] * s.invw, nz = v.c[2] * s.invw;
	const float sx = (nx * 0.5f + 0.5f) * (float)width, sy = (ny * 0.5f + 0.5f) * (float)height;
	s.X = (int32_t)floorf(sx * 256.0f + 0.5f);
	s.Y = (int32_t)floorf(sy * 256.0f + 0.5f);
	s.z = nz * 0.5f + 0.5f;
#pragma unroll
	for (int k = 0; k < kNAttr; k++) s.a[k] = v.a[k];
	return s;
}

__device__ __forceinline__ long long edge(const ScreenV &a, const ScreenV &b, int32_t px, int32_t py) {
	return (long long)(b.X - a.X) * (long long)(py - a.Y) - (long long)(b.Y - a.Y) * (long long)(px - a.X);
}
__device__ __forceinline__ bool owns_edge(const ScreenV &a, const ScreenV &b) {
	const int32_t dx = b.X - a.X, dy = b.Y - a.Y;
	return dy < 0 || (dy == 0 && dx > 0);
}
__device__ __forceinline__ float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }
__device__ __forceinline__ uint32_t to_u8(float v) { return (uint32_t)(clamp01(v) * 255.0f + 0.5f); }

// fragment-shader.glsl:14-24
__device__ __forceinline__ bool fragment(const float *at, uint32_t layer, const RasterParams &p, uint32_t &rgba) {
	const float u = at[1] - floorf(at[1]), v = at[2] - floorf(at[2]);  // Repeat, client.rs:611
	uint32_t iu = (uint32_t)(u * (float)p.ts), iv = (uint32_t)(v * (float)p.ts);
	if (iu >= p.ts) iu = p.ts - 1;
	if (iv >= p.ts) iv = p.ts - 1;
	if (layer >= p.layers) layer = p.layers - 1;
	const float4 t = __ldg(reinterpret_cast<const float4 *>(p.tex) + ((size_t)layer * p.ts + iv) * p.ts + iu);
	if (t.w < 0.5f) return false;  // discard
	const float len = sqrtf(at[3] * at[3] + at[4] * at[4] + at[5] * at[5] + at[6] * at[6]);
	const float fog = clamp01((len - p.fog_far) / p.fog_near);
	const float k = 1.0f - fog;
	const float r = at[0] * t.x * k + 0.5f * fog, g = at[0] * t.y * k + 0.5f * fog, b = at[0] * t.z * k + 0.5f * fog;
	const float a = t.w * k + 1.0f * fog;
	rgba = to_u8(r) | (to_u8(g) << 8) | (to_u8(b) << 16) | (to_u8(a) << 24);
	return true;
}

struct Edges {  // per triangle constants of the inner loop
	bool oa, ob, oc;
	float rfa;  // 1 / (twice the area)
};

// one pixel of the triangle a, b, c (positive area); COVERED: the caller knows the pixel is inside
template <bool COVERED>
__device__ __forceinline__ void pixel(const ScreenV &a, const ScreenV &b, const ScreenV &c, const Edges &e, uint32_t layer,
                                      int32_t x, int32_t y, const RasterParams &p) {
	const int32_t px = x * 256 + 128, py = y * 256 + 128;
	const long long wa = edge(b, c, px, py), wb = edge(c, a, px, py), wc = edge(a, b, px, py);
	if (!COVERED) {
		if (wa < 0 || wb < 0 || wc < 0) return;
		if ((wa == 0 && !e.oa) || (wb == 0 && !e.ob) || (wc == 0 && !e.oc)) return;
	}
	const float la = (float)wa * e.rfa, lb = (float)wb * e.rfa, lc = (float)wc * e.rfa;
	const float depth = la * a.z + lb * b.z + lc * c.z;
	const uint32_t d24 = (uint32_t)(clamp01(depth) * 16777215.0f + 0.5f);
	if (d24 >= 0xFFFFFFu) return;  // IfLess against the cleared 1.0
	const float qa = la * a.invw, qb = lb * b.invw, qc = lc * c.invw;
	const float rq = 1.0f / (qa + qb + qc);
	float at[kNAttr];
#pragma unroll
	for (int k = 0; k < kNAttr; k++) at[k] = (qa * a.a[k] + qb * b.a[k] + qc * c.a[k]) * rq;
	uint32_t rgba;
	if (!fragment(at, layer, p, rgba)) return;
	atomicMin(p.fb + (size_t)y * p.width + x, ((unsigned long long)d24 << 32) | rgba);
}

struct Box {
	int32_t x0, x1, y0, y1;
};

// floor(n / d) for d > 0, saturated at +-2^20 (the caller clamps to a box within +-8192): a float estimate made
// exact with the integer products -- a 64-bit division costs some hundred instructions, and this runs three
// times per row of every listed triangle.  |n|, d < 2^52: the estimate is off by less than one.
__device__ __forceinline__ long long floor_div(long long n, long long d) {
	const float qf = floorf((float)n / (float)d);
	if (qf >= 1048576.0f) return 1 << 20;
	if (qf <= -1048576.0f) return -(1 << 20);
	long long q = (long long)qf;
	while (q * d > n) q--;
	while ((q + 1) * d <= n) q++;
	return q;
}
// The pixels x of row y with edge(u, w, centre) >= 0 (> 0 unless `owns`), intersected with [x0, x1]: the same
// integers the per-pixel test accepts, solved for x.  edge = C - dy * 256 * x with C as below.
__device__ __forceinline__ void clip_span(const ScreenV &u, const ScreenV &w, bool owns, int32_t y, int32_t &x0, int32_t &x1) {
	const long long dx = w.X - u.X, dy = w.Y - u.Y;
	const long long C = dx * (long long)(y * 256 + 128 - u.Y) - dy * (long long)(128 - u.X);
	const long long slack = owns ? 0 : 1;
	if (dy > 0) {          // dy * 256 * x <= C - slack
		const long long hi = floor_div(C - slack, dy * 256);
		if (hi < x1) x1 = (int32_t)max(hi, (long long)x0 - 1);
	} else if (dy < 0) {   // -dy * 256 * x >= slack - C
		const long long lo = -floor_div(C - slack, -dy * 256);  // ceil((slack - C) / (-dy * 256))
		if (lo > x0) x0 = (int32_t)min(lo, (long long)x1 + 1);
	} else if (C - slack < 0) {
		x1 = x0 - 1;
	}
}
__device__ __forceinline__ Box bbox(const ScreenV &a, const ScreenV &b, const ScreenV &c, uint32_t width, uint32_t height) {
	const int32_t minX = min(a.X, min(b.X, c.X)), maxX = max(a.X, max(b.X, c.X));
	const int32_t minY = min(a.Y, min(b.Y, c.Y)), maxY = max(a.Y, max(b.Y, c.Y));
	Box r;
	r.x0 = max((minX - 128 + 255) >> 8, 0);
	r.x1 = min((maxX - 128) >> 8, (int32_t)width - 1);
	r.y0 = max((minY - 128 + 255) >> 8, 0);
	r.y1 = min((maxY - 128) >> 8, (int32_t)height - 1);
	return r;
}

// One set-up triangle p0, p1, p2 in draw order: cull, then rasterise here or queue it.
__device__ void emit_triangle(const ScreenV &p0, const ScreenV &p1, const ScreenV &p2, uint32_t layer, const RasterParams &p) {
	const long long area = edge(p0, p1, p2.X, p2.Y);
	if (area >= 0) return;  // counter-clockwise (or empty): culled, client.rs:667
	const ScreenV &a = p0, &b = p2, &c = p1;  // positive-area orientation
	const Box bx = bbox(a, b, c, p.width, p.height);
	if (bx.x1 < bx.x0 || bx.y1 < bx.y0) return;
	const long long npx = (long long)(bx.x1 - bx.x0 + 1) * (bx.y1 - bx.y0 + 1);
	Edges e;
	e.oa = owns_edge(b, c); e.ob = owns_edge(c, a); e.oc = owns_edge(a, b);
	e.rfa = 1.0f / (float)(-area);
	if (npx > kSmallBox) {
		const int32_t tx0 = bx.x0 >> kTileShift, tx1 = bx.x1 >> kTileShift, ty0 = bx.y0 >> kTileShift, ty1 = bx.y1 >> kTileShift;
		const uint32_t ntiles = (uint32_t)(tx1 - tx0 + 1) * (uint32_t)(ty1 - ty0 + 1);
		const uint32_t slot = atomicAdd(p.big_count, 1u);
		if (slot < kBigCap && bx.y1 - bx.y0 < kMidRows && bx.x1 - bx.x0 < kMidCols) {
			const uint32_t at = atomicAdd(p.big_count + 2, 1u);
			if (at < kMidCap) {
				BigTri &t = p.big[slot];
				t.v[0] = a; t.v[1] = b; t.v[2] = c;
				t.layer = layer;
				p.mid[at] = slot;
				return;
			}
		}
		const uint32_t base = slot < kBigCap ? atomicAdd(p.big_count + 1, ntiles) : kTileCap;
		if (base < kTileCap && ntiles <= kTileCap - base) {
			BigTri &t = p.big[slot];
			t.v[0] = a; t.v[1] = b; t.v[2] = c;
			t.layer = layer;
			uint32_t k = base;
			for (int32_t ty = ty0; ty <= ty1; ty++)
				for (int32_t tx = tx0; tx <= tx1; tx++, k++) {
					// the pixel centres of the tile span [lo, hi]; wholly outside one edge -> nothing to draw there
					const int32_t lx = (tx << kTileShift) * 256 + 128, hx = lx + ((1 << kTileShift) - 1) * 256;
					const int32_t ly = (ty << kTileShift) * 256 + 128, hy = ly + ((1 << kTileShift) - 1) * 256;
					bool keep = true;
#pragma unroll
					for (int q = 0; q < 3; q++) {
						const ScreenV &u = q == 0 ? b : q == 1 ? c : a, &w = q == 0 ? c : q == 1 ? a : b;
						const long long e0 = edge(u, w, lx, ly), e1 = edge(u, w, hx, ly), e2 = edge(u, w, lx, hy), e3 = edge(u, w, hx, hy);
						keep = keep && !(e0 < 0 && e1 < 0 && e2 < 0 && e3 < 0);
					}
					p.tiles[k] = make_uint2(keep ? slot : kNoTri, (uint32_t)tx | ((uint32_t)ty << 16));
				}
			return;
		}
		// lists full: this thread does the work itself; what it had reserved of the tile list stays empty
		for (uint32_t k = base; k < kTileCap && k - base < ntiles; k++) p.tiles[k] = make_uint2(kNoTri, 0u);
	}
	for (int32_t y = bx.y0; y <= bx.y1; y++)
		for (int32_t x = bx.x0; x <= bx.x1; x++) pixel<false>(a, b, c, e, layer, x, y, p);
}

__global__ void __launch_bounds__(kTriThreads) raster_tris_kernel(const RasterParams p) {
	for (uint64_t i = (uint64_t)blockIdx.x * kTriThreads + threadIdx.x; i < p.n_tris; i += (uint64_t)gridDim.x * kTriThreads) {
		ClipV poly[kMaxPoly];
		bool inside = true;  // all three corners inside all six planes: nothing to clip
		float lo[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
		bool out_all[6] = {true, true, true, true, true, true};
#pragma unroll
		for (int k = 0; k < 3; k++) {
			const float4 *s = reinterpret_cast<const float4 *>(p.verts + 3 * i + k);
			const float4 c = __ldg(s), vp = __ldg(s + 1), r = __ldg(s + 2);
			poly[k].c[0] = c.x; poly[k].c[1] = c.y; poly[k].c[2] = c.z; poly[k].c[3] = c.w;
			poly[k].a[0] = r.x; poly[k].a[1] = r.y; poly[k].a[2] = r.z;
			poly[k].a[3] = vp.x; poly[k].a[4] = vp.y; poly[k].a[5] = vp.z; poly[k].a[6] = vp.w;
			if (k == 2) lo[0] = r.w;  // vtex_ind of the provoking (last) vertex, bits kept as they are
#pragma unroll
			for (int q = 0; q < 6; q++) {
				const bool in = plane_dist(poly[k], q) >= 0.0f;
				inside = inside && in;
				out_all[q] = out_all[q] && !in;
			}
		}
		const uint32_t layer = __float_as_uint(lo[0]);
		// outside one plane with all three corners: clipping would leave nothing
		if (out_all[0] || out_all[1] || out_all[2] || out_all[3] || out_all[4] || out_all[5]) continue;
		int m = 3;
		if (!inside) {
			ClipV tmp[kMaxPoly];
			m = clip_poly(poly, tmp, 3);
			if (m < 3) continue;
		}
		const ScreenV s0 = to_screen(poly[0], p.width, p.height);
		ScreenV sp = to_screen(poly[1], p.width, p.height);
		for (int k = 2; k < m; k++) {
			const ScreenV sn = to_screen(poly[k], p.width, p.height);
			emit_triangle(s0, sp, sn, layer, p);
			sp = sn;
		}
	}
}

constexpr int kMidWarps = 4;
__global__ void __launch_bounds__(kMidWarps * 32) raster_mid_kernel(const RasterParams p) {
	__shared__ BigTri ts[kMidWarps];
	__shared__ int32_t span_x0[kMidWarps][32];
	__shared__ uint32_t incl[kMidWarps][32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned full = 0xffffffffu;
	const uint32_t n = min(p.big_count[2], kMidCap);
	for (uint32_t i = blockIdx.x * kMidWarps + warp; i < n; i += gridDim.x * kMidWarps) {
		const uint32_t *src = reinterpret_cast<const uint32_t *>(p.big + __ldg(p.mid + i));
		__syncwarp();
		for (int k = lane; k < (int)(sizeof(BigTri) / 4); k += 32) reinterpret_cast<uint32_t *>(&ts[warp])[k] = src[k];
		__syncwarp();
		const BigTri &t = ts[warp];
		const ScreenV &a = t.v[0], &b = t.v[1], &c = t.v[2];
		const Box bx = bbox(a, b, c, p.width, p.height);
		Edges e;
		e.oa = owns_edge(b, c); e.ob = owns_edge(c, a); e.oc = owns_edge(a, b);
		e.rfa = 1.0f / (float)edge(a, b, c.X, c.Y);
		// lane <-> row
		const int32_t y = bx.y0 + lane;
		int32_t x0 = bx.x0, x1 = y <= bx.y1 ? bx.x1 : bx.x0 - 1;
		if (x1 >= x0) clip_span(b, c, e.oa, y, x0, x1);
		if (x1 >= x0) clip_span(c, a, e.ob, y, x0, x1);
		if (x1 >= x0) clip_span(a, b, e.oc, y, x0, x1);
		uint32_t inc = x1 >= x0 ? (uint32_t)(x1 - x0 + 1) : 0u;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t o = __shfl_up_sync(full, inc, d);
			if (lane >= d) inc += o;
		}
		span_x0[warp][lane] = x0;
		incl[warp][lane] = inc;
		const uint32_t total = __shfl_sync(full, inc, 31);
		__syncwarp();
		// lane <-> covered pixel
		for (uint32_t k = lane; k < total; k += 32) {
			int r = 0;
#pragma unroll
			for (int step = 16; step > 0; step >>= 1)
				if (incl[warp][r + step - 1] <= k) r += step;
			const uint32_t before = r == 0 ? 0u : incl[warp][r - 1];
			pixel<true>(a, b, c, e, t.layer, span_x0[warp][r] + (int32_t)(k - before), bx.y0 + r, p);
		}
	}
}

__global__ void __launch_bounds__(kBigThreads) raster_big_kernel(const RasterParams p) {
	__shared__ BigTri t;
	constexpr int kRows = 1 << kTileShift;
	__shared__ int32_t span_x0[kRows];
	__shared__ uint32_t incl[kRows];  // covered pixels of rows 0..r (rows 0..31) resp. 32..r (rows 32..63)
	const uint32_t n = min(p.big_count[1], kTileCap);
	for (uint32_t i = blockIdx.x; i < n; i += gridDim.x) {
		const uint2 ent = __ldg(p.tiles + i);
		if (ent.x == kNoTri) continue;
		__syncthreads();
		if (threadIdx.x < sizeof(BigTri) / 4) reinterpret_cast<uint32_t *>(&t)[threadIdx.x] = reinterpret_cast<const uint32_t *>(p.big + ent.x)[threadIdx.x];
		__syncthreads();
		const ScreenV &a = t.v[0], &b = t.v[1], &c = t.v[2];
		Box bx = bbox(a, b, c, p.width, p.height);
		const int32_t tx = (int32_t)(ent.y & 0xffffu) << kTileShift, ty = (int32_t)(ent.y >> 16) << kTileShift;
		bx.x0 = max(bx.x0, tx); bx.x1 = min(bx.x1, tx + (1 << kTileShift) - 1);
		bx.y0 = max(bx.y0, ty); bx.y1 = min(bx.y1, ty + (1 << kTileShift) - 1);
		Edges e;
		e.oa = owns_edge(b, c); e.ob = owns_edge(c, a); e.oc = owns_edge(a, b);
		e.rfa = 1.0f / (float)edge(a, b, c.X, c.Y);
		// thread <-> row of the tile: the covered span of the row, exactly
		if (threadIdx.x < kRows) {
			const int32_t y = bx.y0 + (int32_t)threadIdx.x;
			int32_t x0 = bx.x0, x1 = y <= bx.y1 ? bx.x1 : bx.x0 - 1;
			if (x1 >= x0) clip_span(b, c, e.oa, y, x0, x1);
			if (x1 >= x0) clip_span(c, a, e.ob, y, x0, x1);
			if (x1 >= x0) clip_span(a, b, e.oc, y, x0, x1);
			span_x0[threadIdx.x] = x0;
			// exclusive prefix of the span lengths over the 64 rows (two warps)
			const uint32_t len = x1 >= x0 ? (uint32_t)(x1 - x0 + 1) : 0u;
			uint32_t inc = len;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
				if ((int)(threadIdx.x & 31) >= d) inc += o;
			}
			incl[threadIdx.x] = inc;
		}
		__syncthreads();
		const uint32_t half = incl[31], total = half + incl[kRows - 1];
		// thread <-> covered pixel; its row = the first one of its half whose inclusive count exceeds the index
		for (uint32_t k = threadIdx.x; k < total; k += kBigThreads) {
			const int base = k >= half ? 32 : 0;
			const uint32_t kk = k - (base ? half : 0u);
			int r = base;
#pragma unroll
			for (int step = 16; step > 0; step >>= 1)
				if (incl[r + step - 1] <= kk) r += step;
			const uint32_t before = r == base ? 0u : incl[r - 1];
			pixel<true>(a, b, c, e, t.layer, span_x0[r] + (int32_t)(kk - before), bx.y0 + r, p);
		}
	}
}

__global__ void clear_kernel(unsigned long long *fb, size_t n, uint32_t *big_count) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) fb[i] = ~0ull;
	if (i == 0) big_count[0] = big_count[1] = big_count[2] = 0;
}
__global__ void resolve_kernel(const unsigned long long *fb, size_t n, uint32_t clear, uint32_t *image) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) image[i] = fb[i] == ~0ull ? clear : (uint32_t)fb[i];
}

uint32_t host_u8(float v) {
	volatile float c = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
	volatile float s = c * 255.0f;
	s = s + 0.5f;
	return (uint32_t)s;
}

}  // namespace

extern "C" {

int mm_set_textures(mm_ctx *c, const float *rgba, uint32_t layers, uint32_t size) {
	if (!c || !rgba || !layers || !size || size > 4096) return MM_ERR_INVALID;
	MM_CUDA(c, cudaSetDevice(c->device));
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	cudaFree(c->d_tex);
	c->d_tex = nullptr;
	c->tex_layers = c->tex_size = 0;
	const size_t bytes = (size_t)layers * size * size * 16;
	MM_CUDA(c, cudaMalloc(&c->d_tex, bytes));
	MM_CUDA(c, cudaMemcpy(c->d_tex, rgba, bytes, cudaMemcpyHostToDevice));
	c->tex_layers = layers;
	c->tex_size = size;
	return MM_OK;
}

int mm_raster_frame(mm_ctx *c, const mm_shaded_vertex *d_verts, uint64_t n_vertices, const mm_frame *f, uint32_t *d_image,
                    float *kernel_ms) {
	if (!c || !f || !d_image || (n_vertices && !d_verts)) return MM_ERR_INVALID;
	if (!f->width || !f->height || f->width > 8192 || f->height > 8192) return fail(c, MM_ERR_INVALID, "frame size must be 1..8192");
	if (n_vertices % 3) return fail(c, MM_ERR_INVALID, "a triangle list has a multiple of 3 vertices");
	if (!c->d_tex) return fail(c, MM_ERR_NO_TABLE, "mm_set_textures has not been called");
	MM_CUDA(c, cudaSetDevice(c->device));
	const size_t npx = (size_t)f->width * f->height;
	if (npx > c->fb_pixels) {
		MM_CUDA(c, cudaStreamSynchronize(c->stream));
		cudaFree(c->d_fb);
		c->d_fb = nullptr;
		c->fb_pixels = 0;
		MM_CUDA(c, cudaMalloc(&c->d_fb, npx * 8));
		c->fb_pixels = npx;
	}
	if (!c->d_big) {
		MM_CUDA(c, cudaMalloc(&c->d_big, sizeof(BigTri) * (size_t)kBigCap + sizeof(uint2) * (size_t)kTileCap + 4 * (size_t)kMidCap + 256));
		MM_CUDA(c, cudaMalloc(&c->d_big_count, 256));
	}
	if (!c->ev2) {
		MM_CUDA(c, cudaEventCreate(&c->ev2));
		MM_CUDA(c, cudaEventCreate(&c->ev3));
	}
	RasterParams p;
	p.verts = d_verts;
	p.n_tris = n_vertices / 3;
	p.width = f->width; p.height = f->height;
	p.fog_near = f->fog_near; p.fog_far = f->fog_far;
	p.tex = c->d_tex; p.layers = c->tex_layers; p.ts = c->tex_size;
	p.fb = (unsigned long long *)c->d_fb;
	p.big = (BigTri *)c->d_big;
	p.tiles = (uint2 *)(c->d_big + sizeof(BigTri) * (size_t)kBigCap);
	p.mid = (uint32_t *)(c->d_big + sizeof(BigTri) * (size_t)kBigCap + sizeof(uint2) * (size_t)kTileCap);
	p.big_count = (uint32_t *)c->d_big_count;
	const uint32_t clear = host_u8(f->clear[0]) | (host_u8(f->clear[1]) << 8) | (host_u8(f->clear[2]) << 16) | (host_u8(f->clear[3]) << 24);
	const unsigned pxblocks = (unsigned)((npx + 255) / 256);
	MM_CUDA(c, cudaEventRecord(c->ev2, c->stream));
	clear_kernel<<<pxblocks, 256, 0, c->stream>>>(p.fb, npx, p.big_count);
	if (p.n_tris) {
		const uint64_t want = (p.n_tris + kTriThreads - 1) / kTriThreads;
		const unsigned grid = (unsigned)std::min<uint64_t>(want, (uint64_t)c->sm_count * 64);
		raster_tris_kernel<<<grid, kTriThreads, 0, c->stream>>>(p);
		raster_mid_kernel<<<c->sm_count * 16, kMidWarps * 32, 0, c->stream>>>(p);
		raster_big_kernel<<<c->sm_count * 32, kBigThreads, 0, c->stream>>>(p);
	}
	resolve_kernel<<<pxblocks, 256, 0, c->stream>>>(p.fb, npx, clear, d_image);
	MM_CUDA(c, cudaGetLastError());
	MM_CUDA(c, cudaEventRecord(c->ev3, c->stream));
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	if (kernel_ms) MM_CUDA(c, cudaEventElapsedTime(kernel_ms, c->ev2, c->ev3));
	return MM_OK;
}

}  // extern "C"
