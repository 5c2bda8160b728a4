/* TEST INFRASTRUCTURE -- CPU restatement of what happens to the shaded vertices after the vertex stage
 * (the optional part): primitive assembly, clipping, back-face culling,
 * rasterisation, depth test and the fragment shader, into a headless RGBA8 image.  Never linked into
 * the product library.
 *
 *  - draw state: mimas-client/client.rs:661-675 -- triangle lists, depth test IfLess with depth write,
 *    BackfaceCullingMode::CullCounterClockwise, clear colour (0.05, 0.01, 0.6, 0.0), clear depth 1.0
 *  - fragment shader: mimas-client/fragment-shader.glsl:14-24 (texture array look-up, alpha test at 0.5,
 *    Lambert factor, fog `clamp((length(vposition) - fog_near_far.y) / fog_near_far.x, 0, 1)`,
 *    `mix(colour, (0.5, 0.5, 0.5, 1), fog)`)
 *  - sampler: client.rs:610-612 (wrap Repeat, magnify Nearest); textures are 16 x 16 float RGBA layers
 *    (mimas-client/assets.rs:194-211)
 *
 * PARITY UNPINNED, and by construction not pixel-exact to any GL driver: OpenGL leaves rasterisation
 * (sub-pixel snapping, tie rules), depth quantisation and float evaluation order to the implementation,
 * GL is unavailable on this box and the reference has no images to compare with.  This file fixes one
 * choice of all of them; the CUDA kernels (mimas_b200/csrc/mm_raster.cu) are held bit-exact to it.
 * Deliberate simplifications, the same on both sides:
 *  - minification samples level 0 with the nearest texel as well (glium's default minify filter is
 *    LinearMipmapLinear over auto-generated mipmaps);
 *  - no sRGB decode/encode (texels and the image are taken as they are);
 *  - a fragment that survives the alpha test is written unblended (the reference blends with
 *    Blend::alpha_blending in HashMap order, client.rs:668,699-708; with texel alpha 0 or 1, which is
 *    what the game's textures have, and without fog that is the same thing);
 *  - of two fragments with the same 24-bit depth the one with the smaller packed colour wins (GL: the
 *    one drawn first, in an unspecified order).
 *
 * Conventions: window coordinates as in GL (origin bottom-left, pixel centres at +0.5), vertices
 * snapped to 1/256 pixel, a pixel centre exactly on an edge belongs to the triangle whose edge runs
 * downwards (or horizontally to the right) in the positive-area orientation.  Every float operation is
 * written out singly rounded, left to right (compile with -ffp-contract=off). */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct ra_shaded {  /* = mm_shaded_vertex */
	float clip[4];
	float vposition[4];
	float vlamb;
	float vtex_pos[2];
	uint32_t vtex_ind;
} ra_shaded;

typedef struct ra_frame {  /* = mm_frame */
	uint32_t width, height;
	float fog_near, fog_far;
	float clear[4];
} ra_frame;

#define NATTR 7      /* vlamb, u, v, vposition.xyzw */
#define GUARD 16.0f  /* guard band: clip x, y at +-GUARD * w */
#define MAXV 12

typedef struct {
	float c[4];
	float a[NATTR];
} cv;

static float plane_dist(const cv *v, int p) {
	switch (p) {
	case 0: return v->c[2] + v->c[3];          /* near:  z >= -w */
	case 1: return v->c[3] - v->c[2];          /* far:   z <=  w */
	case 2: return v->c[0] + GUARD * v->c[3];
	case 3: return GUARD * v->c[3] - v->c[0];
	case 4: return v->c[1] + GUARD * v->c[3];
	default: return GUARD * v->c[3] - v->c[1];
	}
}

/* the point where the edge from `in` (distance di >= 0) to `out` (distance dq < 0) meets the plane */
static cv cut(const cv *in, const cv *out, float di, float dq) {
	const float t = di / (di - dq);
	cv r;
	for (int k = 0; k < 4; k++) r.c[k] = in->c[k] + t * (out->c[k] - in->c[k]);
	for (int k = 0; k < NATTR; k++) r.a[k] = in->a[k] + t * (out->a[k] - in->a[k]);
	return r;
}

static int clip_poly(cv *poly, int n) {
	cv tmp[MAXV];
	for (int p = 0; p < 6 && n; p++) {
		int m = 0;
		for (int i = 0; i < n; i++) {
			const cv *a = &poly[i], *b = &poly[(i + 1) % n];
			const float da = plane_dist(a, p), db = plane_dist(b, p);
			if (da >= 0.0f) {
				tmp[m++] = *a;
				if (db < 0.0f) tmp[m++] = cut(a, b, da, db);
			} else if (db >= 0.0f) {
				tmp[m++] = cut(b, a, db, da);
			}
		}
		n = m;
		memcpy(poly, tmp, sizeof(cv) * (size_t)n);
	}
	return n;
}

typedef struct {
	int32_t X, Y;  /* window coordinates, 1/256 pixel */
	float z, invw;
	const float *a;
} sv;

static sv to_screen(const cv *v, const ra_frame *f) {
	sv s;
	s.invw = 1.0f / v->c[3];
	const float nx = v->c[0] * s.invw, ny = v->c[1] * s.invw, nz = v->c[2] * s.invw;
	const float sx = (nx * 0.5f + 0.5f) * (float)f->width, sy = (ny * 0.5f + 0.5f) * (float)f->height;
	s.X = (int32_t)floorf(sx * 256.0f + 0.5f);
	s.Y = (int32_t)floorf(sy * 256.0f + 0.5f);
	s.z = nz * 0.5f + 0.5f;
	s.a = v->a;
	return s;
}

static int64_t edge(const sv *a, const sv *b, int64_t px, int64_t py) {
	return (int64_t)(b->X - a->X) * (py - a->Y) - (int64_t)(b->Y - a->Y) * (px - a->X);
}
/* a pixel centre exactly on the edge a -> b is inside iff the edge runs down, or level and to the right */
static int owns_edge(const sv *a, const sv *b) {
	const int32_t dx = b->X - a->X, dy = b->Y - a->Y;
	return dy < 0 || (dy == 0 && dx > 0);
}

static float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }
static uint32_t to_u8(float v) { return (uint32_t)(clamp01(v) * 255.0f + 0.5f); }

/* fragment-shader.glsl:14-24; returns 0 if the fragment is discarded */
static int fragment(const float *at, uint32_t layer, const ra_frame *f, const float *tex, uint32_t layers, uint32_t ts,
		uint32_t *rgba) {
	const float u = at[1] - floorf(at[1]), v = at[2] - floorf(at[2]);  /* Repeat */
	uint32_t iu = (uint32_t)(u * (float)ts), iv = (uint32_t)(v * (float)ts);
	if (iu >= ts) iu = ts - 1;
	if (iv >= ts) iv = ts - 1;
	if (layer >= layers) layer = layers - 1;
	const float *t = tex + (((size_t)layer * ts + iv) * ts + iu) * 4;
	if (t[3] < 0.5f) return 0;
	const float len = sqrtf(at[3] * at[3] + at[4] * at[4] + at[5] * at[5] + at[6] * at[6]);
	const float fog = clamp01((len - f->fog_far) / f->fog_near);
	const float k = 1.0f - fog;
	const float r = at[0] * t[0] * k + 0.5f * fog, g = at[0] * t[1] * k + 0.5f * fog, b = at[0] * t[2] * k + 0.5f * fog;
	const float a = t[3] * k + 1.0f * fog;
	*rgba = to_u8(r) | (to_u8(g) << 8) | (to_u8(b) << 16) | (to_u8(a) << 24);
	return 1;
}

static void raster_tri(const sv *p0, const sv *p1, const sv *p2, uint32_t layer, const ra_frame *f, const float *tex,
		uint32_t layers, uint32_t ts, uint64_t *fb) {
	const int64_t area = edge(p0, p1, p2->X, p2->Y);
	if (area >= 0) return;  /* counter-clockwise (or empty): culled, client.rs:667 */
	/* positive-area orientation: p0, p2, p1 */
	const sv *a = p0, *b = p2, *c = p1;
	const int64_t A = -area;
	int32_t minX = a->X, maxX = a->X, minY = a->Y, maxY = a->Y;
	if (b->X < minX) minX = b->X;
	if (c->X < minX) minX = c->X;
	if (b->X > maxX) maxX = b->X;
	if (c->X > maxX) maxX = c->X;
	if (b->Y < minY) minY = b->Y;
	if (c->Y < minY) minY = c->Y;
	if (b->Y > maxY) maxY = b->Y;
	if (c->Y > maxY) maxY = c->Y;
	int32_t x0 = (minX - 128 + 255) >> 8, x1 = (maxX - 128) >> 8, y0 = (minY - 128 + 255) >> 8, y1 = (maxY - 128) >> 8;
	if (x0 < 0) x0 = 0;
	if (y0 < 0) y0 = 0;
	if (x1 > (int32_t)f->width - 1) x1 = (int32_t)f->width - 1;
	if (y1 > (int32_t)f->height - 1) y1 = (int32_t)f->height - 1;
	const int oa = owns_edge(b, c), ob = owns_edge(c, a), oc = owns_edge(a, b);
	const float rfa = 1.0f / (float)A;
	for (int32_t y = y0; y <= y1; y++)
		for (int32_t x = x0; x <= x1; x++) {
			const int64_t px = (int64_t)x * 256 + 128, py = (int64_t)y * 256 + 128;
			const int64_t wa = edge(b, c, px, py), wb = edge(c, a, px, py), wc = edge(a, b, px, py);
			if (wa < 0 || wb < 0 || wc < 0) continue;
			if ((wa == 0 && !oa) || (wb == 0 && !ob) || (wc == 0 && !oc)) continue;
			const float la = (float)wa * rfa, lb = (float)wb * rfa, lc = (float)wc * rfa;
			const float depth = la * a->z + lb * b->z + lc * c->z;
			const uint32_t d24 = (uint32_t)(clamp01(depth) * 16777215.0f + 0.5f);
			if (d24 >= 0xFFFFFFu) continue;  /* IfLess against the cleared 1.0 */
			const float qa = la * a->invw, qb = lb * b->invw, qc = lc * c->invw;
			const float rq = 1.0f / (qa + qb + qc);
			float at[NATTR];
			for (int k = 0; k < NATTR; k++) at[k] = (qa * a->a[k] + qb * b->a[k] + qc * c->a[k]) * rq;
			uint32_t rgba;
			if (!fragment(at, layer, f, tex, layers, ts, &rgba)) continue;
			const uint64_t key = ((uint64_t)d24 << 32) | rgba;
			uint64_t *dst = fb + (size_t)y * f->width + x;
			if (key < *dst) *dst = key;
		}
}

/* n vertices (triangle list, n % 3 == 0) -> image[height][width] RGBA8 packed R | G<<8 | B<<16 | A<<24, row 0 at the
 * bottom; `keys` (width*height u64 scratch, may be NULL) receives depth24 << 32 | colour per pixel. */
int ra_raster_frame(const ra_shaded *v, uint64_t n, const ra_frame *f, const float *tex, uint32_t layers, uint32_t ts,
		uint32_t *image, uint64_t *keys_out) {
	const size_t npx = (size_t)f->width * f->height;
	uint64_t *fb = keys_out;
	uint64_t *own = NULL;
	if (!fb) {
		own = fb = (uint64_t *)malloc(npx * 8);
		if (!fb) return 1;
	}
	memset(fb, 0xff, npx * 8);
	for (uint64_t i = 0; i + 2 < n; i += 3) {
		cv poly[MAXV];
		for (int k = 0; k < 3; k++) {
			const ra_shaded *s = v + i + k;
			memcpy(poly[k].c, s->clip, 16);
			poly[k].a[0] = s->vlamb;
			poly[k].a[1] = s->vtex_pos[0];
			poly[k].a[2] = s->vtex_pos[1];
			memcpy(poly[k].a + 3, s->vposition, 16);
		}
		const uint32_t layer = v[i + 2].vtex_ind;  /* flat: the last vertex provokes */
		const int m = clip_poly(poly, 3);
		if (m < 3) continue;
		sv s[MAXV];
		for (int k = 0; k < m; k++) s[k] = to_screen(&poly[k], f);
		for (int k = 1; k + 1 < m; k++) raster_tri(&s[0], &s[k], &s[k + 1], layer, f, tex, layers, ts, fb);
	}
	const uint32_t clear = to_u8(f->clear[0]) | (to_u8(f->clear[1]) << 8) | (to_u8(f->clear[2]) << 16) | (to_u8(f->clear[3]) << 24);
	for (size_t p = 0; p < npx; p++) image[p] = fb[p] == ~(uint64_t)0 ? clear : (uint32_t)fb[p];
	free(own);
	return 0;
}
