/*
 * meshgen_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C99) of the mimas chunk mesher, used ONLY as the parity
 * checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs.  Nothing under mimas_b200/ may link, load or call this file.
 *
 * PARITY UNPINNED: the reference (est31/mimas, Rust) ships no tests, golden vectors or
 * fixtures for mimas-meshgen, and no rustc/cargo exists on this box, so this
 * restatement is pinned only by (a) a line-by-line reading of the source cited below,
 * (b) an independent table-driven derivation of known-answer vectors
 * (tests/golden/make_kats.py) and (c) an independent Python transliteration
 * (oracle/meshgen_oracle.py).  All three agree bit-for-bit (tests/test_oracle_*.py).
 *
 * Every function cites the reference file:line it follows (paths relative to the
 * reference checkout, /root/reference in the build container).
 *
 * Build: see oracle/Makefile (gcc -O3 -ffp-contract=off; no -march: the .so is built in one container and run on another host).  FP contraction
 * is off because Rust never fuses a*b+c; all arithmetic here is IEEE binary32 add/sub/
 * mul/neg in round-to-nearest-even, which x86-64 SSE scalar code reproduces exactly.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <time.h>

#define CHUNKSIZE 16 /* mimas-common/map.rs:11 */

/* mimas-meshgen/lib.rs:18-24 `Vertex`, laid out #[repr(C)] in declared order
 * (SURVEY.md §8b): tex_ind:u16 @0, 2 B zero padding @2, tex_pos @4, position @12,
 * normal @24; size 36, align 4. */
typedef struct {
	uint16_t tex_ind;
	uint16_t pad0;
	float tex_pos[2];
	float position[3];
	float normal[3];
} mo_vertex;

/* mimas-meshgen/lib.rs:44-49 `BlockTextureIds` */
typedef struct {
	uint16_t id_sides, id_top, id_bottom;
} mo_bti;

/* mimas-meshgen/lib.rs:73-79 `TextureIdCache` flattened.  `len` is
 * block_texture_ids.len() == mesh_draw_styles.len() == block_params.len().
 * style: 0 = None, 1 = Some(Blocky), 2 = Some(Crossed)   (lib.rs:29-32, :89-94) */
typedef struct {
	mo_bti fallback_id;
	uint32_t len;
	uint8_t style[256];
	mo_bti blocky[256];
	uint16_t crossed[256];
} mo_cache;

/* Growable vector following Rust's Vec<T>: starts empty without allocation
 * (Vec::new(), lib.rs:233,412) and grows by RawVec::grow_amortized:
 * cap = max(2*cap, len+1, 4).  Kept so that the CPU baseline pays the same
 * reallocation pattern as the reference. */
typedef struct {
	mo_vertex *data;
	size_t len, cap;
} mo_vec;

static void vec_push(mo_vec *v, mo_vertex x) {
	if (v->len == v->cap) {
		size_t nc = v->cap * 2;
		if (nc < v->len + 1) nc = v->len + 1;
		if (nc < 4) nc = 4;
		v->data = (mo_vertex *)realloc(v->data, nc * sizeof(mo_vertex));
		v->cap = nc;
	}
	v->data[v->len++] = x;
}

/* mimas-meshgen/lib.rs:123-131 `zsig` */
static float zsig(float v) {
	if (v > 0.0f) return 1.0f;
	if (v < 0.0f) return -1.0f;
	return 0.0f;
}

static mo_vertex mkv(uint16_t tex, float u, float v, float px, float py, float pz,
		float nx, float ny, float nz) {
	mo_vertex r;
	r.tex_ind = tex;
	r.pad0 = 0;
	r.tex_pos[0] = u; r.tex_pos[1] = v;
	r.position[0] = px; r.position[1] = py; r.position[2] = pz;
	/* sign![..] = [zsig(..), zsig(..), zsig(..)], lib.rs:133-137 */
	r.normal[0] = zsig(nx); r.normal[1] = zsig(ny); r.normal[2] = zsig(nz);
	return r;
}

/* mimas-meshgen/lib.rs:139-149 `rpush_face!` -- expression for expression. */
static void rpush_face(mo_vec *r, float x, float y, float z,
		float xsd, float ysd, float yd, float zd, uint16_t tex_ind) {
	vec_push(r, mkv(tex_ind, 0.0f, 0.0f, x, y, z, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, xsd + ysd, 0.0f, x + xsd, y + ysd, z, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, 0.0f, yd + zd, x, y + yd, z + zd, xsd, ysd + yd, zd));

	vec_push(r, mkv(tex_ind, xsd + ysd, 0.0f, x + xsd, y + ysd, z, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, xsd + ysd, yd + zd, x + xsd, y + yd + ysd, z + zd, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, 0.0f, yd + zd, x, y + yd, z + zd, xsd, ysd + yd, zd));
}

/* mimas-meshgen/lib.rs:150-160 `rpush_face_rev!` */
static void rpush_face_rev(mo_vec *r, float x, float y, float z,
		float xsd, float ysd, float yd, float zd, uint16_t tex_ind) {
	vec_push(r, mkv(tex_ind, 0.0f, yd + zd, x, y + yd, z + zd, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, xsd + ysd, 0.0f, x + xsd, y + ysd, z, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, 0.0f, 0.0f, x, y, z, -xsd, -ysd - yd, -zd));

	vec_push(r, mkv(tex_ind, 0.0f, yd + zd, x, y + yd, z + zd, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, xsd + ysd, yd + zd, x + xsd, y + yd + ysd, z + zd, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, xsd + ysd, 0.0f, x + xsd, y + ysd, z, -xsd, -ysd - yd, -zd));
}

/* mimas-meshgen/lib.rs:162-181 `rpush_face_bidi!` */
static void rpush_face_bidi(mo_vec *r, float x, float y, float z,
		float xsd, float ysd, float yd, float zd,
		float xstd, float ystd, float ytd, float ztd, uint16_t tex_ind) {
	vec_push(r, mkv(tex_ind, 0.0f, 0.0f, x, y, z, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, xstd + ystd, 0.0f, x + xsd, y + ysd, z, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, 0.0f, ytd + ztd, x, y + yd, z + zd, xsd, ysd + yd, zd));

	vec_push(r, mkv(tex_ind, xstd + ystd, 0.0f, x + xsd, y + ysd, z, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, xstd + ystd, ytd + ztd, x + xsd, y + yd + ysd, z + zd, xsd, ysd + yd, zd));
	vec_push(r, mkv(tex_ind, 0.0f, ytd + ztd, x, y + yd, z + zd, xsd, ysd + yd, zd));

	/* Reverse face */
	vec_push(r, mkv(tex_ind, 0.0f, ytd + ztd, x, y + yd, z + zd, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, xstd + ystd, 0.0f, x + xsd, y + ysd, z, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, 0.0f, 0.0f, x, y, z, -xsd, -ysd - yd, -zd));

	vec_push(r, mkv(tex_ind, 0.0f, ytd + ztd, x, y + yd, z + zd, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, xstd + ystd, ytd + ztd, x + xsd, y + yd + ysd, z + zd, -xsd, -ysd - yd, -zd));
	vec_push(r, mkv(tex_ind, xstd + ystd, 0.0f, x + xsd, y + ysd, z, -xsd, -ysd - yd, -zd));
}

/* mimas-common/map.rs:83-86 `MapChunkData::get_blk` */
static uint8_t get_blk(const uint8_t *chunk, const int64_t p[3]) {
	return chunk[p[0] * CHUNKSIZE * CHUNKSIZE + p[1] * CHUNKSIZE + p[2]];
}

/* mimas-meshgen/lib.rs:105-109 `get_bl_tex_ids`: known id -> stored Option
 * (None for undrawn and for crossed ids, lib.rs:92-94); unknown id -> Some(fallback). */
static int get_bl_tex_ids(const mo_cache *c, uint8_t id, mo_bti *out) {
	if ((uint32_t)id < c->len) {
		if (c->style[id] == 1) { *out = c->blocky[id]; return 1; }
		return 0;
	}
	*out = c->fallback_id;
	return 1;
}

/* mimas-meshgen/lib.rs:273-283 `blocked`.
 * `halo` == NULL is the reference: a neighbour outside the chunk never blocks (lib.rs:277-280).
 * A non-NULL halo is the OPTIONAL extension of this repo (no reference counterpart):
 * 6 slabs of 256 block ids, the layer of each face-neighbour chunk that touches this chunk, in the
 * order x-, x+, y-, y+, z-, z+; x slabs indexed y*16+z, y slabs x*16+z, z slabs x*16+y.  A blocky
 * cell there blocks the face exactly as a cell inside the chunk would; nothing else changes. */
static int blocked(const uint8_t *chunk, const uint8_t *halo, const int64_t o[3], const int64_t pos[3],
		const mo_cache *cache) {
	int64_t p[3] = { pos[0] + o[0], pos[1] + o[1], pos[2] + o[2] };
	mo_bti t;
	for (int i = 0; i < 3; i++)
		if (p[i] < 0 || p[i] >= CHUNKSIZE) {
			if (!halo) return 0; /* outside => not blocked */
			const int64_t u = i == 0 ? p[1] : p[0], v = i == 2 ? p[1] : p[2];  /* the two in-range coordinates */
			return get_bl_tex_ids(cache, halo[(2 * i + (p[i] < 0 ? 0 : 1)) * 256 + u * CHUNKSIZE + v], &t);
		}
	return get_bl_tex_ids(cache, get_blk(chunk, p), &t);
}

/* mimas-meshgen/lib.rs:284-293 `get_tex_ind` */
static int get_tex_ind(const uint8_t *chunk, const uint8_t *halo, const int64_t pos[3], const int64_t offsets[3],
		const mo_cache *cache, mo_bti *out) {
	mo_bti t;
	int some = get_bl_tex_ids(cache, get_blk(chunk, pos), &t);
	if (some && blocked(chunk, halo, offsets, pos, cache)) return 0;
	if (some) *out = t;
	return some;
}

/* mimas-meshgen/lib.rs:235-272 `Walker<TextureId>` */
typedef struct {
	int has_last;
	float last_v;
	uint16_t l_item;
} walker;

/* One of the six passes of lib.rs:316-410, described by data instead of closures:
 *   inner_axis : which coordinate `cinner` drives in f(c1,c2,cinner) (lib.rs:318,334,...)
 *   face       : 0 = id_sides, 1 = id_top, 2 = id_bottom          (lib.rs:319,335,...)
 *   rev        : rpush_face_rev! instead of rpush_face!
 *   the emit closure bodies are in emit_pass() below. */
typedef struct {
	int id;          /* 1..6 in source order */
	int inner_axis;  /* 0 = x, 1 = y */
	int face;
	int64_t offsets[3];
} pass_desc;

static const pass_desc PASSES[6] = {
	{ 1, 1, 2, { 0, 0, -1 } }, /* lib.rs:317-330  X-Y face (unify over y), id_bottom */
	{ 2, 0, 0, { 0, -1, 0 } }, /* lib.rs:333-346  X-Z face (unify over x), id_sides  */
	{ 3, 1, 0, { -1, 0, 0 } }, /* lib.rs:349-362  Y-Z face (unify over y), id_sides  */
	{ 4, 1, 1, { 0, 0, 1 } },  /* lib.rs:365-378  X-Y face (z+1),          id_top    */
	{ 5, 0, 0, { 0, 1, 0 } },  /* lib.rs:381-394  X-Z face (y+1),          id_sides  */
	{ 6, 1, 0, { 1, 0, 0 } },  /* lib.rs:397-410  Y-Z face (x+1),          id_sides  */
};

/* The `emit` closures handed to walker.next in each pass (lib.rs:324-327, 340-343,
 * 356-359, 372-375, 388-391, 404-407).  `pos` is the position of the cell at which
 * the run ENDED (the closure captures `pos` of the current g() call), of which only
 * the two non-inner coordinates are used. */
static void emit_pass(int pass, mo_vec *r, const int64_t pos[3], uint16_t tx,
		float last, float len) {
	const float siz = 1.0f; /* lib.rs:314 */
	float x = (float)pos[0], y = (float)pos[1], z = (float)pos[2];
	switch (pass) {
	case 1: rpush_face(r, x, last, z, siz, 0.0f, len, 0.0f, tx); break;
	case 2: rpush_face_rev(r, last, y, z, len, 0.0f, 0.0f, siz, tx); break;
	case 3: rpush_face(r, x, last, z, 0.0f, len, 0.0f, siz, tx); break;
	case 4: rpush_face_rev(r, x, last, z + siz, siz, 0.0f, len, 0.0f, tx); break;
	case 5: rpush_face(r, last, y + siz, z, len, 0.0f, 0.0f, siz, tx); break;
	case 6: rpush_face_rev(r, x + siz, last, z, 0.0f, len, 0.0f, siz, tx); break;
	}
}

/* mimas-meshgen/lib.rs:244-271 `Walker::next` with the pass's emit closure inlined. */
static void walker_next(walker *w, float v, int has_item, uint16_t item,
		int pass, mo_vec *r, const int64_t pos[3]) {
	if (!has_item && w->has_last) {
		/* Some mesh ends here. Emit it. */
		float vlen = v - w->last_v;
		emit_pass(pass, r, pos, w->l_item, w->last_v, vlen);
		w->has_last = 0;
	} else if (has_item && w->has_last) {
		if (item != w->l_item) {
			/* Item changed. Emit the old item. */
			float vlen = v - w->last_v;
			emit_pass(pass, r, pos, w->l_item, w->last_v, vlen);
			w->has_last = 1; w->last_v = v; w->l_item = item;
		}
	} else if (has_item && !w->has_last) {
		w->has_last = 1; w->last_v = v; w->l_item = item;
	}
}

/* mimas-meshgen/lib.rs:294-313 `walk_for_all_blocks` + the `g` closures
 * (lib.rs:322-328 etc.: pos = offs + rel_pos; walker.next(pos.<inner> as f32, ...)). */
static void walk_for_all_blocks(const pass_desc *p, const int64_t offs[3],
		const uint8_t *chunk, const uint8_t *halo, const mo_cache *cache, mo_vec *r) {
	for (int64_t c1 = 0; c1 < CHUNKSIZE; c1++) {
		for (int64_t c2 = 0; c2 < CHUNKSIZE; c2++) {
			walker w = { 0, 0.0f, 0 };
			for (int64_t cinner = 0; cinner <= CHUNKSIZE; cinner++) {
				/* f(c1, c2, cinner): lib.rs:318 (c1,cinner,c2) / :334 (cinner,c1,c2) */
				int64_t rel[3];
				if (p->inner_axis == 1) { rel[0] = c1; rel[1] = cinner; rel[2] = c2; }
				else { rel[0] = cinner; rel[1] = c1; rel[2] = c2; }
				int has = 0;
				uint16_t tex_ind = 0;
				if (cinner < CHUNKSIZE) {
					mo_bti t;
					has = get_tex_ind(chunk, halo, rel, p->offsets, cache, &t);
					if (has) tex_ind = p->face == 0 ? t.id_sides : p->face == 1 ? t.id_top : t.id_bottom;
				} /* else: the sentinel g(&mut walker, None, f(c1,c2,CHUNKSIZE)), lib.rs:309-310 */
				int64_t pos[3] = { offs[0] + rel[0], offs[1] + rel[1], offs[2] + rel[2] };
				float v = (float)pos[p->inner_axis];
				walker_next(&w, v, has, tex_ind, p->id, r, pos);
			}
		}
	}
}

/* mimas-meshgen/lib.rs:231-446 `mesh_for_chunk` */
static void mesh_for_chunk_h(const int64_t offs[3], const uint8_t *chunk, const uint8_t *halo, const mo_cache *cache,
		mo_vec *r, mo_vec *rt) {
	for (int i = 0; i < 6; i++)
		walk_for_all_blocks(&PASSES[i], offs, chunk, halo, cache, r);

	/* Crossed nodes, lib.rs:414-440 */
	const float siz = 1.0f;
	for (int64_t x = 0; x < CHUNKSIZE; x++) {
		for (int64_t y = 0; y < CHUNKSIZE; y++) {
			for (int64_t z = 0; z < CHUNKSIZE; z++) {
				int64_t rel[3] = { x, y, z };
				uint8_t blk = get_blk(chunk, rel);
				/* get_mesh_draw_style, lib.rs:110-113: unknown id -> None */
				if (!((uint32_t)blk < cache->len && cache->style[blk] == 2)) continue;
				uint16_t tx = cache->crossed[blk];
				int64_t pos[3] = { offs[0] + x, offs[1] + y, offs[2] + z };

				float sq = siz * 1.41421356237309504880168872420969808f * 0.5f; /* lib.rs:428 */
				float sqh = sq * 0.5f;                                          /* lib.rs:429 */
				float fx = (float)pos[0] + 0.5f * siz;                          /* lib.rs:431 */
				float fy = (float)pos[1] + 0.5f * siz;
				float fz = (float)pos[2];
				float tsiz = siz * 0.5f;                                        /* lib.rs:433 */
				/* X-Z, lib.rs:435 */
				rpush_face_bidi(rt, fx - sqh, fy - sqh, fz, sq, sq, 0.0f, siz * 0.95f,
					tsiz, tsiz, 0.0f, siz * 0.95f, tx);
				/* Y-Z, lib.rs:437 */
				rpush_face_bidi(rt, fx + sqh, fy - sqh, fz, -sq, sq, 0.0f, siz * 0.95f,
					tsiz, tsiz, 0.0f, siz * 0.95f, tx);
			}
		}
	}
}

/* mimas-meshgen/lib.rs:183-219 `push_block` (un-merged cube emitter).  `blocked_mask`
 * bit i answers the i-th blocked([..]) query in source order:
 * [0,0,-1], [0,-1,0], [-1,0,0], [0,0,1], [0,1,0], [1,0,0]. */
static void push_block(mo_vec *r, float x, float y, float z, mo_bti ids, float siz,
		unsigned blocked_mask) {
	if (!(blocked_mask & 1))  rpush_face(r, x, y, z, siz, 0.0f, siz, 0.0f, ids.id_bottom);
	if (!(blocked_mask & 2))  rpush_face_rev(r, x, y, z, siz, 0.0f, 0.0f, siz, ids.id_sides);
	if (!(blocked_mask & 4))  rpush_face(r, x, y, z, 0.0f, siz, 0.0f, siz, ids.id_sides);
	if (!(blocked_mask & 8))  rpush_face_rev(r, x, y, z + siz, siz, 0.0f, siz, 0.0f, ids.id_top);
	if (!(blocked_mask & 16)) rpush_face(r, x, y + siz, z, siz, 0.0f, 0.0f, siz, ids.id_sides);
	if (!(blocked_mask & 32)) rpush_face_rev(r, x + siz, y, z, 0.0f, siz, 0.0f, siz, ids.id_sides);
}

static void mesh_for_chunk(const int64_t offs[3], const uint8_t *chunk, const mo_cache *cache,
		mo_vec *r, mo_vec *rt) {
	mesh_for_chunk_h(offs, chunk, 0, cache, r, rt);
}

/* ------------------------------------------------------------------------------------
 * C entry points (ctypes-facing).  All "mo_" = meshgen oracle.
 * ---------------------------------------------------------------------------------- */

/* Mesh one chunk.  Returns two malloc'd vertex arrays (caller frees with mo_free). */
void mo_mesh_for_chunk(const int64_t offs[3], const uint8_t *chunk, const mo_cache *cache,
		mo_vertex **intransparent, uint64_t *n_intransparent,
		mo_vertex **transparent, uint64_t *n_transparent) {
	mo_vec r = { 0, 0, 0 }, rt = { 0, 0, 0 };
	mesh_for_chunk(offs, chunk, cache, &r, &rt);
	*intransparent = r.data; *n_intransparent = r.len;
	*transparent = rt.data; *n_transparent = rt.len;
}

void mo_free(void *p) { free(p); }

void mo_push_block(float x, float y, float z, const mo_bti *ids, float siz,
		uint32_t blocked_mask, mo_vertex **out, uint64_t *n_out) {
	mo_vec r = { 0, 0, 0 };
	push_block(&r, x, y, z, *ids, siz, blocked_mask);
	*out = r.data; *n_out = r.len;
}

/* Count pass over a batch: per chunk vertex counts only (cheap way to size buffers). */
void mo_count_batch_halo(const uint8_t *chunks, const int64_t *offs, uint64_t n, const mo_cache *cache,
		uint32_t *n_intransparent, uint32_t *n_transparent, const uint8_t *halo);
void mo_count_batch(const uint8_t *chunks, const int64_t *offs, uint64_t n, const mo_cache *cache,
		uint32_t *n_intransparent, uint32_t *n_transparent) {
	mo_count_batch_halo(chunks, offs, n, cache, n_intransparent, n_transparent, 0);
}
void mo_count_batch_halo(const uint8_t *chunks, const int64_t *offs, uint64_t n, const mo_cache *cache,
		uint32_t *n_intransparent, uint32_t *n_transparent, const uint8_t *halo) {
	for (uint64_t i = 0; i < n; i++) {
		mo_vec r = { 0, 0, 0 }, rt = { 0, 0, 0 };
		mesh_for_chunk_h(offs + 3 * i, chunks + 4096 * i, halo ? halo + 1536 * i : 0, cache, &r, &rt);
		n_intransparent[i] = (uint32_t)r.len; n_transparent[i] = (uint32_t)rt.len;
		free(r.data); free(rt.data);
	}
}

/* Mesh a batch into caller-provided packed streams.  first_* are exclusive prefix sums
 * of the per-chunk counts (as produced from mo_count_batch); the streams hold the
 * chunks' vertices back to back in chunk order -- the same packing the CUDA library
 * produces, so a whole batch can be compared with one memcmp. */
typedef struct {
	const uint8_t *chunks; const int64_t *offs; const mo_cache *cache; const uint8_t *halo;
	uint64_t n;
	uint64_t *next; /* shared work counter: the threads draw blocks of MO_GRAIN chunks until the batch is done */
	mo_vertex *intr; mo_vertex *trans;
	const uint64_t *first_intr; const uint64_t *first_trans;
	uint64_t verts; /* out */
} batch_job;

#define MO_GRAIN 4u
static void *batch_worker(void *arg) {
	batch_job *j = (batch_job *)arg;
	uint64_t verts = 0;
	for (;;) {
		const uint64_t begin = __atomic_fetch_add(j->next, MO_GRAIN, __ATOMIC_RELAXED);
		if (begin >= j->n) break;
		const uint64_t end = begin + MO_GRAIN < j->n ? begin + MO_GRAIN : j->n;
		for (uint64_t i = begin; i < end; i++) {
			/* fresh, empty vectors per chunk, as in the reference (lib.rs:233,412) */
			mo_vec r = { 0, 0, 0 }, rt = { 0, 0, 0 };
			mesh_for_chunk_h(j->offs + 3 * i, j->chunks + 4096 * i, j->halo ? j->halo + 1536 * i : 0, j->cache, &r, &rt);
			if (j->intr && r.len) memcpy(j->intr + j->first_intr[i], r.data, r.len * sizeof(mo_vertex));
			if (j->trans && rt.len) memcpy(j->trans + j->first_trans[i], rt.data, rt.len * sizeof(mo_vertex));
			verts += r.len + rt.len;
			free(r.data); free(rt.data);
		}
	}
	j->verts = verts;
	return 0;
}

/* Runs the batch on `threads` OS threads (chunks are independent, lib.rs:277-280); the threads share one
 * atomic work counter, so that a region whose chunks differ a lot in cost (a quarter of a terrain region
 * is all air) keeps every core busy to the end.  intr/trans may be NULL (mesh and discard: the CPU
 * baseline's timing mode -- the meshes are still fully built in per-chunk vectors).
 * Returns total vertices produced; *seconds receives the wall time of the parallel
 * region (CLOCK_MONOTONIC). */
uint64_t mo_mesh_batch_halo(const uint8_t *chunks, const int64_t *offs, uint64_t n, const mo_cache *cache,
		mo_vertex *intr, const uint64_t *first_intr,
		mo_vertex *trans, const uint64_t *first_trans,
		uint32_t threads, double *seconds, const uint8_t *halo);
uint64_t mo_mesh_batch(const uint8_t *chunks, const int64_t *offs, uint64_t n, const mo_cache *cache,
		mo_vertex *intr, const uint64_t *first_intr,
		mo_vertex *trans, const uint64_t *first_trans,
		uint32_t threads, double *seconds) {
	return mo_mesh_batch_halo(chunks, offs, n, cache, intr, first_intr, trans, first_trans, threads, seconds, 0);
}
/* the same with the optional halo (n * 1536 bytes) */
uint64_t mo_mesh_batch_halo(const uint8_t *chunks, const int64_t *offs, uint64_t n, const mo_cache *cache,
		mo_vertex *intr, const uint64_t *first_intr,
		mo_vertex *trans, const uint64_t *first_trans,
		uint32_t threads, double *seconds, const uint8_t *halo) {
	if (threads < 1) threads = 1;
	if (threads > 1024) threads = 1024;
	if ((uint64_t)threads > n && n > 0) threads = (uint32_t)n;
	batch_job *jobs = (batch_job *)calloc(threads, sizeof(batch_job));
	pthread_t *tids = (pthread_t *)calloc(threads, sizeof(pthread_t));
	struct timespec t0, t1;
	uint64_t next = 0;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (uint32_t t = 0; t < threads; t++) {
		batch_job *j = &jobs[t];
		j->chunks = chunks; j->offs = offs; j->cache = cache; j->halo = halo;
		j->n = n; j->next = &next;
		j->intr = intr; j->trans = trans; j->first_intr = first_intr; j->first_trans = first_trans;
		if (threads == 1) batch_worker(j);
		else pthread_create(&tids[t], 0, batch_worker, j);
	}
	uint64_t verts = 0;
	for (uint32_t t = 0; t < threads; t++) {
		if (threads > 1) pthread_join(tids[t], 0);
		verts += jobs[t].verts;
	}
	clock_gettime(CLOCK_MONOTONIC, &t1);
	if (seconds) *seconds = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
	free(jobs); free(tids);
	return verts;
}

uint32_t mo_sizeof_vertex(void) { return (uint32_t)sizeof(mo_vertex); }
uint32_t mo_sizeof_cache(void) { return (uint32_t)sizeof(mo_cache); }
