/*
 * mimas_meshgen_cuda.h -- C ABI of libmimas_meshgen_cuda.so
 *
 * B200 (sm_100a) CUDA replacement for the hot path of est31/mimas'
 * `mimas-meshgen` crate.  Every entry point names the reference interface it
 * stands in for (paths relative to the reference checkout).  The Rust crate
 * keeps its public API (`mesh_for_chunk`, `TextureIdCache`, `push_block`) and
 * binds these symbols through `extern "C"` -- see INTEGRATION.md for the stub.
 *
 * Conventions: plain pointers and sizes only; every function returns an
 * mm_status (0 = OK) and never unwinds; an mm_ctx is single-threaded (one per
 * calling thread / GPU), like the reference's dedicated meshgen thread
 * (mimas-client/client.rs:259-267).  There is no CPU fallback: without a CUDA
 * device mm_ctx_create fails with MM_ERR_NO_DEVICE.
 */
#ifndef MIMAS_MESHGEN_CUDA_H
#define MIMAS_MESHGEN_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MM_ABI_VERSION 1u
#define MM_CHUNKSIZE 16u           /* mimas-common/map.rs:11  CHUNKSIZE            */
#define MM_CHUNK_BYTES 4096u       /* mimas-common/map.rs:56-61 [MapBlock; 16^3]   */
#define MM_HALO_BYTES 1536u        /* optional halo of one chunk: 6 slabs of 256 block ids (extension) */
#define MM_TABLE_ENTRIES 256u      /* MapBlock is a u8 id, mimas-common/map.rs:14  */

typedef enum mm_status {
	MM_OK = 0,
	MM_ERR_INVALID = 1,     /* bad argument                                          */
	MM_ERR_NO_DEVICE = 2,   /* no usable CUDA device (there is no CPU fallback)      */
	MM_ERR_CUDA = 3,        /* CUDA runtime error, see mm_last_error_string          */
	MM_ERR_OOM = 4,         /* device or pinned host allocation failed               */
	MM_ERR_NO_TABLE = 5,    /* mm_set_block_table has not been called                */
	MM_ERR_NO_BATCH = 6,    /* no finished batch to read back                        */
	MM_ERR_UNSUPPORTED = 7
} mm_status;

/* mimas-meshgen/lib.rs:18-24 `Vertex`, fixed as #[repr(C)] in declared order:
 * tex_ind @0, two zero padding bytes @2, tex_pos @4, position @12, normal @24;
 * size 36, align 4.  The device stream and every host copy use exactly this. */
typedef struct mm_vertex {
	uint16_t tex_ind;
	uint16_t pad0;      /* always written as 0 so that streams can be byte-compared */
	float tex_pos[2];
	float position[3];
	float normal[3];
} mm_vertex;

/* One row of the flattened `TextureIdCache` (mimas-meshgen/lib.rs:73-79): what
 * get_bl_tex_ids (lib.rs:105-109) and get_mesh_draw_style (lib.rs:110-113)
 * answer for block id i.  Ids >= block_params.len() must be filled by the caller
 * as MM_KIND_BLOCKY with the fallback ids (lib.rs:106-108) -- mm_table_from_cache
 * does that. */
#define MM_KIND_NONE 0u     /* not drawn: None in both vectors                       */
#define MM_KIND_BLOCKY 1u   /* Some(MeshDrawStyle::Blocky(BlockTextureIds))           */
#define MM_KIND_CROSSED 2u  /* Some(MeshDrawStyle::Crossed(TextureId))                */
typedef struct mm_block_entry {
	uint16_t id_sides;   /* BlockTextureIds::id_sides,  lib.rs:46 */
	uint16_t id_top;     /* BlockTextureIds::id_top,    lib.rs:47 */
	uint16_t id_bottom;  /* BlockTextureIds::id_bottom, lib.rs:48 */
	uint16_t id_crossed; /* MeshDrawStyle::Crossed(id), lib.rs:31 */
	uint8_t kind;        /* MM_KIND_*                             */
	uint8_t reserved[3];
} mm_block_entry;

/* Where one chunk's vertices live inside a packed stream (unit: vertices). */
typedef struct mm_range {
	uint64_t first_vertex;
	uint32_t vertex_count;
	uint32_t reserved;
} mm_range;

/* Result of a batch: the two `ChunkMesh` vectors (lib.rs:226-229) of every chunk,
 * packed back to back in chunk order, device resident.  Pointers stay valid until
 * the next mm_mesh_batch* / mm_reserve / mm_ctx_destroy on the same context. */
typedef struct mm_batch_out {
	const mm_vertex *d_intransparent; /* device */
	const mm_vertex *d_transparent;   /* device */
	const mm_range *d_ranges_i;       /* device, n_chunks entries */
	const mm_range *d_ranges_t;       /* device, n_chunks entries */
	uint64_t n_intransparent;         /* total vertices in d_intransparent */
	uint64_t n_transparent;           /* total vertices in d_transparent   */
	uint32_t n_chunks;
	uint32_t launches;                /* kernels launched for this batch (incl. re-runs after growth) */
	float kernel_ms;                  /* CUDA-event time of this batch's kernels (mm_mesh_batch); 0 for mm_mesh_batch_async batches */
	uint32_t reserved;
} mm_batch_out;

typedef struct mm_ctx mm_ctx;

uint32_t mm_abi_version(void);

/* Context = one GPU + one stream + device block table + output arenas.  Replaces
 * the state the reference's meshgen thread owns (its TextureIdCache clone,
 * mimas-client/client.rs:258-261). */
int mm_ctx_create(int device, mm_ctx **out);
int mm_ctx_destroy(mm_ctx *ctx);
const char *mm_last_error_string(const mm_ctx *ctx);

/* Run on a caller-owned CUDA stream (cudaStream_t passed as void*), e.g. the
 * framework's current stream; NULL restores the context's own stream. */
int mm_ctx_set_stream(mm_ctx *ctx, void *cuda_stream);

/* Device-resident block table.  Replaces the TextureIdCache handed to
 * mesh_for_chunk (lib.rs:231-232, built by TextureIdCache::from_hdl, lib.rs:87-104). */
int mm_set_block_table(mm_ctx *ctx, const mm_block_entry table[MM_TABLE_ENTRIES]);

/* Helper for bindings: flatten a TextureIdCache given as its three vectors
 * (lib.rs:75-77): fallback ids, and per known id a kind + ids row.  Fills ids
 * >= n_known with the blocky fallback (lib.rs:105-109). */
int mm_table_from_cache(const uint16_t fallback_sides_top_bottom[3],
                        const mm_block_entry *known, uint32_t n_known,
                        mm_block_entry out_table[MM_TABLE_ENTRIES]);

/* Pre-size the arenas (vertices).  Optional: batches grow them on demand and
 * re-run, which costs one extra kernel the first time a larger batch is seen. */
int mm_reserve(mm_ctx *ctx, uint32_t max_chunks, uint64_t intransparent_vertices,
               uint64_t transparent_vertices);

/* mesh_for_chunk (lib.rs:231-446) over a batch of n independent chunks.
 *   chunks : n * 4096 block ids, index x*256 + y*16 + z (mimas-common/map.rs:83-86);
 *            host or device memory (detected)
 *   offs   : n * 3 int64 chunk origins in blocks (`offs :Vector3<isize>`, lib.rs:231)
 *            host or device memory (detected)
 *   halo   : NULL = the reference's semantics: out-of-chunk neighbours never block
 *            (lib.rs:277-280).  Non-NULL = an EXTENSION with no counterpart in the reference
 *            (parity: oracle/meshgen_oracle.c `blocked` with halo slabs): n * MM_HALO_BYTES
 *            block ids, per chunk the six layers of the face-neighbour chunks that touch it, in
 *            the order x-, x+, y-, y+, z-, z+, 256 ids each; the x slabs indexed y*16 + z, the
 *            y slabs x*16 + z, the z slabs x*16 + y.  A face against a blocky cell of a slab is
 *            culled like one against a blocky cell of the chunk; runs, textures and the crossed
 *            pass are untouched.  Host or device memory (detected); device memory must be
 *            16-byte aligned.  mm_gather_halos builds the slabs for a resident region.
 * Synchronous: returns after the kernel finished and totals are known. */
int mm_mesh_batch(mm_ctx *ctx, const uint8_t *chunks, const int64_t *offs,
                  const uint8_t *halo, uint32_t n, mm_batch_out *out);

/* Stream-ordered variant for device-resident inputs (16-byte aligned chunks, 8-byte aligned offs):
 * enqueues the mesher kernels and returns immediately.  mm_batch_finish waits for the most recent
 * batch, grows + re-runs it if its arenas were too small, and reports totals.  Several _async calls
 * may be queued back to back: consecutive batches overlap on the device.  With the default ring of one
 * output slot only the last one's result is retrievable; see mm_ring_configure / mm_mesh_batch_submit. */
int mm_mesh_batch_async(mm_ctx *ctx, const uint8_t *d_chunks, const int64_t *d_offs, uint32_t n);

/* Output ring (new; the batch submission path for mapgen chunks of the north star): `slots` (1..8) library-owned
 * sets of arenas.  Every batch -- mm_mesh_batch_submit, and equally mm_mesh_batch / _async -- takes the next
 * ticket t = 1, 2, ... and writes its packed streams and range records to slot t % slots, so up to `slots`
 * batches are retrievable at once while later ones are being meshed.  mm_batch_wait(t) blocks until batch t has
 * landed (the device signals completion per batch through mapped host memory: no stream synchronisation, the
 * batches behind it keep running), grows + re-runs it if its slot was too small (its inputs must still be
 * resident), and fills `out`; the pointers stay valid until ticket t + slots is submitted, after which
 * mm_batch_wait(t) answers MM_ERR_NO_BATCH.  mm_reserve sizes every slot.  Reconfiguring drops outstanding tickets. */
int mm_ring_configure(mm_ctx *ctx, uint32_t slots);
int mm_mesh_batch_submit(mm_ctx *ctx, const uint8_t *d_chunks, const int64_t *d_offs, const uint8_t *d_halo, uint32_t n,
                         uint64_t *ticket);
/* k batches (repeated `repeat` >= 1 times round robin) enqueued by one call, so that a host language with slow
 * foreign calls is not in the 30 us launch cadence; *first_ticket receives the first batch's ticket. */
int mm_mesh_batches_submit(mm_ctx *ctx, const uint8_t *const *d_chunks, const int64_t *const *d_offs, const uint32_t *n,
                           uint32_t k, uint32_t repeat, uint64_t *first_ticket);
int mm_batch_wait(mm_ctx *ctx, uint64_t ticket, mm_batch_out *out);
int mm_batch_poll(mm_ctx *ctx, uint64_t ticket, int *done);
/* mm_copy_out for the batch of `ticket` (waits for it first), on the context's consumer stream. */
int mm_copy_out_ticket(mm_ctx *ctx, uint64_t ticket, mm_vertex *intransparent, mm_vertex *transparent,
                       mm_range *ranges_i, mm_range *ranges_t);
/* The same with halo slabs (device memory, 16-byte aligned; NULL = mm_mesh_batch_async). */
int mm_mesh_batch_halo_async(mm_ctx *ctx, const uint8_t *d_chunks, const int64_t *d_offs, const uint8_t *d_halo,
                             uint32_t n);
/* Halo slabs of a device-resident box of nx * ny * nz chunks laid out like mm_gen_terrain_* lays
 * them out (chunk i at (i % nx, (i / nx) % ny, i / (nx*ny))): every chunk's six slabs from its
 * neighbours in the box, `outside_id` beyond the box.  d_halo: nx*ny*nz * MM_HALO_BYTES,
 * 16-byte aligned.  Stream-ordered. */
int mm_gather_halos(mm_ctx *ctx, const uint8_t *d_chunks, uint32_t nx, uint32_t ny, uint32_t nz, uint8_t outside_id,
                    uint8_t *d_halo);
int mm_batch_finish(mm_ctx *ctx, mm_batch_out *out);

/* Host-`Vec` path: copy the last finished batch to caller memory sized from
 * mm_batch_out (any pointer may be NULL to skip that part). */
int mm_copy_out(mm_ctx *ctx, mm_vertex *intransparent, mm_vertex *transparent,
                mm_range *ranges_i, mm_range *ranges_t);

/* Host in, host out in one call (H2D + kernel + D2H, pipelined in sub-batches): the
 * batch form of the crate API `chunk in, Vec<Vertex> out`.  The library owns the
 * returned host arrays (pinned) until the next call on this context. */
typedef struct mm_host_mesh {
	const mm_vertex *intransparent;
	const mm_vertex *transparent;
	const mm_range *ranges_i;
	const mm_range *ranges_t;
	uint64_t n_intransparent, n_transparent;
	uint32_t n_chunks;
	uint32_t path;      /* MM_PATH_*: which host path the call took */
} mm_host_mesh;
#define MM_PATH_MAPPED 1u     /* small batch: kernels read and write mapped pinned host memory, no copies */
#define MM_PATH_PIPELINED 2u  /* large batch cut into pieces: H2D, kernels and D2H overlap */
#define MM_PATH_PLAIN 3u      /* H2D, kernels, D2H one after the other (first large call, or after an arena overflow) */
int mm_mesh_batch_host(mm_ctx *ctx, const uint8_t *chunks, const int64_t *offs, uint32_t n,
                       mm_host_mesh *out);

/* Single-chunk convenience with exactly mesh_for_chunk's shape (lib.rs:231-232):
 * pointers are library-owned pinned host memory, valid until the next call. */
int mm_mesh_for_chunk(mm_ctx *ctx, const int64_t offs[3], const uint8_t chunk[MM_CHUNK_BYTES],
                      const mm_vertex **intransparent, uint64_t *n_intransparent,
                      const mm_vertex **transparent, uint64_t *n_transparent);

/* push_block (lib.rs:183-219): un-merged cube emitter for UI meshes.  `blocked_mask`
 * bit i is the answer of the i-th blocked([..]) query in source order
 * ([0,0,-1],[0,-1,0],[-1,0,0],[0,0,1],[0,1,0],[1,0,0]).  n cubes in, device-built,
 * copied to `out` (capacity 36 vertices per cube); *n_out receives the count. */
typedef struct mm_cube {
	float pos[3];
	float siz;
	uint16_t id_sides, id_top, id_bottom;
	uint16_t blocked_mask;
} mm_cube;
int mm_push_blocks(mm_ctx *ctx, const mm_cube *cubes, uint32_t n, mm_vertex *out, uint64_t *n_out);

/* Synthetic mapgen look-alike (benchmark input; NOT a replica of mimas-server
 * mapgen, whose noise crates are not available -- see DESIGN.md).  Chunk i of the
 * region has chunk coordinates (cx0 + i % nx, cy0 + (i / nx) % ny, cz0 + i / (nx*ny));
 * offs = 16 * chunk coordinate.  Host and device versions produce identical bytes. */
int mm_gen_terrain_host(uint64_t seed, int64_t cx0, int64_t cy0, int64_t cz0,
                        uint32_t nx, uint32_t ny, uint32_t nz, uint8_t *chunks, int64_t *offs);
int mm_gen_terrain_device(mm_ctx *ctx, uint64_t seed, int64_t cx0, int64_t cy0, int64_t cz0,
                          uint32_t nx, uint32_t ny, uint32_t nz, uint8_t *d_chunks, int64_t *d_offs);

/* mimas-server mapgen PHASE ONE on the device (mimas-server/mapgen.rs:64-245 `gen_chunk_phase_one`): the
 * chunks of a box, generated straight into the mesher's input layout (same chunk order as mm_gen_terrain_*), so
 * that server-side mapgen chunks reach the mesher without a host copy.  Phase two (trees, mapgen.rs:357-434) is not
 * part of it: d_tree_counts (NULL or nx*ny*nz words) receives the number of tree spawn points per chunk
 * (mapgen.rs:228).  Ores and plants are the `[[mapgen.ore]]` / `[[mapgen.plant]]` tables of game-params.toml
 * (mimas-common/game_params.rs:134-149, 648-718), the block ids the roles of game_params.rs:79-90.
 * PARITY UNPINNED against the reference (its arithmetic lives in the crates noise 0.7.0, rand_pcg 0.3.1, rand,
 * twox-hash 1.6.1, none of them under the reference checkout; no reference test pins mapgen output): bit-exact to
 * oracle/mapgen_oracle.c, whose PCG and XXH64 parts are pinned by their published vectors.  Synchronous. */
#define MM_MAPGEN_MAX_ORES 8u
#define MM_MAPGEN_MAX_PLANTS 16u
typedef struct mm_mapgen_ore {      /* game_params.rs:134-143 `Ore` */
	uint8_t block;
	uint8_t noise_seed[8];
	uint8_t pcg_seed[8];
	double freq, pcg_chance, limit_a, limit_b;
	int64_t limit_boundary;
} mm_mapgen_ore;
typedef struct mm_mapgen_plant {    /* game_params.rs:145-149 `Plant` */
	uint8_t block;
	uint8_t pcg_seed[8];
	double pcg_limit;
} mm_mapgen_plant;
typedef struct mm_mapgen_params {
	uint8_t air, water, sand, ground, ground_top, stone;   /* BlockRoles, game_params.rs:79-90 */
	uint32_t n_ores, n_plants;
	mm_mapgen_ore ores[MM_MAPGEN_MAX_ORES];
	mm_mapgen_plant plants[MM_MAPGEN_MAX_PLANTS];
} mm_mapgen_params;
int mm_gen_mapgen_device(mm_ctx *ctx, uint64_t seed, const mm_mapgen_params *params, int64_t cx0, int64_t cy0, int64_t cz0,
                         uint32_t nx, uint32_t ny, uint32_t nz, uint8_t *d_chunks, int64_t *d_offs, uint32_t *d_tree_counts);

/* The mesher for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same result as mm_mesh_batch(halo = NULL) for n <= MM_NOALLOC_MAX_CHUNKS chunks, bit for bit, but the caller owns
 * every output and the call allocates nothing (no cudaMalloc / cudaFree / cudaHostAlloc, in any call, first or
 * later): the analysis -> expansion hand-over lives in a static device array of
 * MM_NOALLOC_MAX_CHUNKS x 64 KiB + 2.25 KiB = 16.0 MiB that every GPU which loads the library holds until exit.
 * All pointers are device pointers: chunks (n x 4096, 16-byte aligned), offs (n x 3), d_out_i / d_out_t vertex
 * buffers of cap_*_vertices vertices (16-byte aligned; a capacity of 0 is allowed), d_ranges_i / d_ranges_t (n
 * records each), d_totals (2 words: QUADS in each stream, x 6 = vertices).  The two kernels are enqueued on the
 * context's stream and the call returns; mm_ctx_sync waits.  A stream whose total exceeds its capacity is not
 * written at all (range records and totals still are): compare d_totals with the capacity.
 * n > MM_NOALLOC_MAX_CHUNKS is MM_ERR_INVALID with a message.  The block table must be set (mm_set_block_table,
 * at set-up; that call does allocate).  Every call on every context of a device shares the one scratch array:
 * calls on one stream run in order and are safe; calls that overlap on the GPU (two contexts / streams) race. */
#define MM_NOALLOC_MAX_CHUNKS 256u
int mm_mesh_batch_noalloc(mm_ctx *ctx, const uint8_t *d_chunks, const int64_t *d_offs, uint32_t n, mm_vertex *d_out_i,
                          uint64_t cap_i_vertices, mm_vertex *d_out_t, uint64_t cap_t_vertices, mm_range *d_ranges_i,
                          mm_range *d_ranges_t, uint64_t *d_totals);

/* Plain device memory helpers so that non-CUDA host code (ctypes, Rust) can keep
 * inputs resident without linking the CUDA runtime itself. */
int mm_dev_alloc(mm_ctx *ctx, size_t bytes, void **d_ptr);
int mm_dev_free(mm_ctx *ctx, void *d_ptr);
int mm_dev_upload(mm_ctx *ctx, void *d_dst, const void *h_src, size_t bytes);
int mm_dev_download(mm_ctx *ctx, void *h_dst, const void *d_src, size_t bytes);
int mm_ctx_sync(mm_ctx *ctx);
/* Page-locked host memory: inputs built here are DMA'd directly by mm_mesh_batch_host /
 * mm_mesh_batch (no staging copy); pageable inputs are staged through an internal pinned buffer. */
int mm_host_alloc(mm_ctx *ctx, size_t bytes, void **h_ptr);
int mm_host_free(mm_ctx *ctx, void *h_ptr);

/* Introspection for the bench: SM count and the grid the mesher kernel uses. */
int mm_device_info(mm_ctx *ctx, int *sm_count, int *ctas_per_sm, int *threads_per_cta);
/* Number of mesher kernels this context has launched since creation. */
int mm_launch_count(mm_ctx *ctx, uint64_t *launches);

/* ---- headless consumer of the vertex stream (GL itself is unavailable on this box) ----
 * What the reference does with the meshes every frame, up to the rasteriser: chunk culling
 * (mimas-client/client.rs:679-696: viewing range on the chunk origin, then frustum_query's sphere test on
 * the chunk centre with r = 16*sqrt(3)) and the vertex shader (mimas-client/vertex-shader.glsl:18-27).
 * Matrices are the `[[f32; 4]; 4]` uniforms of client.rs:599-600: column-major, m[4*col + row]. */
typedef struct mm_camera {
	float vmatrix[16];       /* Camera::get_matrix, client.rs:1388-1393 */
	float pmatrix[16];       /* Camera::get_perspective, client.rs:1395-1400 */
	float pos[3];            /* camera.pos, client.rs:677 */
	float viewing_range;     /* config.viewing_range, client.rs:682 */
} mm_camera;
typedef struct mm_shaded_vertex {   /* the outputs of vertex-shader.glsl */
	float clip[4];           /* gl_Position */
	float vposition[4];
	float vlamb;
	float vtex_pos[2];
	uint32_t vtex_ind;
} mm_shaded_vertex;
typedef struct mm_draw_out {
	const mm_shaded_vertex *d_shaded;  /* visible chunks in chunk order: all intransparent meshes, then all transparent (client.rs:699-701) */
	const uint32_t *d_draw_list;       /* indices of the drawn chunks, ascending */
	uint64_t n_shaded_i, n_shaded_t;   /* vertices */
	uint32_t n_drawn, n_chunks;
	float kernel_ms;
	uint32_t reserved;
} mm_draw_out;
/* vertex shader over n device-resident vertices */
int mm_vertex_stage(mm_ctx *ctx, const mm_vertex *d_in, uint64_t n, const mm_camera *cam, mm_shaded_vertex *d_out);
/* client.rs:679-696 over n chunk origins (host or device pointer); h_draw (n entries, host) receives the drawn indices */
int mm_cull_chunks(mm_ctx *ctx, const int64_t *offs, uint32_t n, const mm_camera *cam, uint32_t *h_draw, uint32_t *n_drawn);
/* one frame over the last finished device batch (mm_mesh_batch / mm_batch_finish): cull its chunks, shade the
 * visible ones.  Results stay in library-owned device memory until the next call. */
int mm_draw_batch(mm_ctx *ctx, const mm_camera *cam, mm_draw_out *out);
/* the same over the batch of a ring ticket (waits for it), on the context's consumer stream: batches queued
 * behind it keep meshing while this one is drawn. */
int mm_draw_ticket(mm_ctx *ctx, uint64_t ticket, const mm_camera *cam, mm_draw_out *out);
/* In two halves, so that the host can queue more mesher batches while a frame is shaded: _async waits for the
 * ticket and enqueues the frame on the consumer stream, mm_draw_collect waits for it and reports.  One frame in
 * flight per context. */
int mm_draw_ticket_async(mm_ctx *ctx, uint64_t ticket, const mm_camera *cam);
int mm_draw_collect(mm_ctx *ctx, mm_draw_out *out);

/* What GL does with the shaded vertices, headless: primitive assembly, clipping, back-face culling
 * (CullCounterClockwise), rasterisation, depth test IfLess and the fragment shader
 * (mimas-client/fragment-shader.glsl:14-24) under the draw state of mimas-client/client.rs:661-675, into an
 * RGBA8 image.  Parity to the reference is unpinned (GL leaves the rasterisation rules to the driver); the
 * kernels are bit-exact to oracle/raster_oracle.c, which states the rules chosen and the simplifications
 * (nearest texel at level 0, no sRGB conversion, no blending of surviving fragments). */
typedef struct mm_frame {
	uint32_t width, height;    /* 1..8192 */
	float fog_near, fog_far;   /* uniform fog_near_far, client.rs:618; fog_near > 0 */
	float clear[4];            /* (0.05, 0.01, 0.6, 0.0), client.rs:675 */
} mm_frame;
/* The texture array (mimas-client/assets.rs:194-211): `layers` square images of size x size float RGBA texels,
 * layer = TextureId, row 0 at v = 0.  Host memory; copied. */
int mm_set_textures(mm_ctx *ctx, const float *rgba, uint32_t layers, uint32_t size);
/* n_vertices (a triangle list, device memory, e.g. mm_draw_out.d_shaded) -> d_image[height][width], packed
 * R | G << 8 | B << 16 | A << 24, row 0 at the bottom as in GL (device memory).  Synchronous. */
int mm_raster_frame(mm_ctx *ctx, const mm_shaded_vertex *d_verts, uint64_t n_vertices, const mm_frame *frame,
                    uint32_t *d_image, float *kernel_ms);

/* Development aid, no reference counterpart: device timeline (%globaltimer, ns) of the launch pairs since
 * the previous call, as a ring of `pairs` records of 8 words: analyse {first CTA start, -, last CTA out of
 * work, last CTA end}, expand {first CTA start, first CTA past its dependency wait, -, last CTA end}.
 * Copies the ring recorded so far (as many records as the previous call armed) into `out` (may be NULL),
 * then re-arms with `pairs` records (0 = off). */
int mm_debug_trace(mm_ctx *ctx, uint32_t pairs, uint64_t *out, uint64_t *n_launched);

#ifdef __cplusplus
}
#endif
#endif /* MIMAS_MESHGEN_CUDA_H */
