// mm_internal.h -- context struct shared by the translation units of the library.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdint.h>

#include "../../include/mimas_meshgen_cuda.h"
#include "mesh_core.cuh"

// One output slot of the ring: the arenas a batch's packed vertex streams and range records go to, where its
// totals reach the host, and what is needed to run it again should the arenas turn out too small.
struct mm_slot {
	unsigned char *d_out_i = nullptr, *d_out_t = nullptr;  // arenas (capacities in quads of 216 B)
	uint64_t cap_i = 0, cap_t = 0;
	uint32_t cap_chunks = 0;                               // per-chunk arrays
	mm::Range *d_ranges_i = nullptr, *d_ranges_t = nullptr;
	mm::u64 *h_totals = nullptr;    // pinned + mapped, 2 words: the expand kernel writes the totals straight to the host
	mm::u64 *d_h_totals = nullptr;  // the same words as the device sees them
	// the batch the slot holds
	uint64_t ticket = 0;            // 0 = none yet
	uint64_t wait_ticket = 0;       // the batch is complete once the device has signalled this ticket (h_done)
	bool pending = false, finished = false;
	const unsigned char *last_chunks = nullptr;
	const unsigned char *last_halo = nullptr;
	const long long *last_offs = nullptr;
	uint32_t last_n = 0;
	uint32_t last_launches = 0;
	bool last_timed = false;
	uint64_t tot_i = 0, tot_t = 0;  // quads
	float last_ms = 0.f;
};
constexpr uint32_t kMaxRingSlots = 8;

struct mm_ctx {
	int device = 0;
	int sm_count = 0;
	int ctas_per_sm = 0;
	int analyse_ctas_per_sm = 0;  // CTAs per SM the analyse grid is sized for
	cudaStream_t own_stream = nullptr;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	bool have_table = false;
	mm::TabEntry *d_table = nullptr;

	// output ring (mm_ring_configure; one slot unless asked otherwise): batch with ticket t goes to slot t % ring_slots
	mm_slot ring[kMaxRingSlots];
	uint32_t ring_slots = 1;
	mm_slot *cur = &ring[0];          // slot of the most recent batch: what finish / copy_out / draw_batch refer to
	uint64_t next_ticket = 1;         // tickets start at 1
	uint64_t last_signal_ticket = 0;  // completion signal of the most recent batch that launched kernels (counts on its own)
	uint64_t signalled_upto = 0;      // newest signal some later launch carries (AnalyseParams::h_done)
	mm::u64 *h_host_totals = nullptr, *d_h_host_totals = nullptr;  // pinned + mapped, 2 words: totals of the host paths' launches
	mm::u64 *h_flag = nullptr, *d_h_flag = nullptr;  // pinned + mapped: the fused single-chunk launch of the mapped path says "landed" here
	mm::u64 flag_seq = 0, flag_wanted = 0, flag_armed = 0;
	mm::u64 *h_done = nullptr;        // pinned + mapped: ticket of the last batch whose kernels have completed
	mm::u64 *d_h_done = nullptr;
	cudaStream_t s_draw = nullptr;    // consumers of finished ring slots run here, beside the mesher's stream
	// analysis -> expansion hand-over (mesh_kernel.cuh): per-chunk descriptor slots and counts
	// -- two of each: consecutive launch pairs alternate, so that the analysis of one batch can run
	// beside the expansion of the previous one (programmatic dependent launch)
	uint32_t *d_desc_i = nullptr, *d_desc_t = nullptr;  // 2 x cap_desc_chunks * kDescStrideI / kDescStrideT words
	uint32_t cap_desc_chunks = 0;
	uint32_t *d_cnt = nullptr;        // 2 x { cnt_i[kMaxLaunchChunks], cnt_t[kMaxLaunchChunks] }
	int pair_parity = 0;
	int expand_grid = 0;
	bool use_pdl = true;
	bool fuse_single = true;  // n == 1: analysis and expansion in one launch (MM_NO_FUSED_SINGLE)
	mm::u64 *d_ticket = nullptr;      // [0] ticket counter, [2..3] and [4..5] alternating totals slots
	int totals_slot = 0;              // slot the last launch wrote
	mm::u64 ticket_base = 0;
	unsigned char *d_flags = nullptr; // 256 flag bytes derived from the block table
	bool time_next = false;
	// mm_debug_trace: device timeline of the last trace_pairs launch pairs (8 words per pair, ring)
	mm::u64 *d_trace = nullptr;
	uint32_t trace_pairs = 0;
	uint64_t trace_next = 0;

	// staged inputs for host callers
	unsigned char *d_in_chunks = nullptr;
	long long *d_in_offs = nullptr;
	uint32_t cap_in = 0;
	unsigned char *d_in_halo = nullptr;  // halo mode: staging for halos handed in as host memory
	uint32_t cap_in_halo = 0;
	unsigned char *h_in = nullptr;  // pinned staging: chunks then offs
	size_t h_in_bytes = 0;

	// small-batch path: one mapped pinned block holding inputs, range records and both vertex streams
	unsigned char *h_small = nullptr, *d_small = nullptr;
	bool no_small = false;
	uint32_t mapped_max_chunks = 256, mapped_chunks = 0;  // batches up to this size take the mapped path; capacity of the block
	uint64_t mapped_qi = 0, mapped_qt = 0;               // capacities of the mapped arenas (quads)

	// host results (pinned)
	unsigned char *h_out = nullptr;
	size_t h_out_bytes = 0;
	// pipelined host path (large batches): pieces go in on s_in, come back on s_out, kernels between on `stream`
	cudaStream_t s_in = nullptr, s_out = nullptr;
	cudaEvent_t ev_in[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
	uint32_t pipe_min_chunks = 1024;  // 0: off (MM_PIPE_MIN_CHUNKS)
	double host_qi_per_chunk = -1.0, host_qt_per_chunk = -1.0;  // quads per chunk of the last large host batch

	uint64_t total_launches = 0;
	uint32_t host_path = 0;  // MM_PATH_* of the last mm_mesh_batch_host call

	// headless draw (mm_render.cu): scratch of mm_cull_chunks / mm_draw_batch
	unsigned char *d_cull = nullptr;   // flags, draw list, per-visible-chunk output offsets, block sums
	size_t cull_bytes = 0;
	unsigned char *d_shaded = nullptr;
	size_t shaded_bytes = 0;
	cudaEvent_t ev2 = nullptr, ev3 = nullptr;
	const mm_slot *draw_slot = nullptr;   // the frame in flight (mm_draw_ticket_async .. mm_draw_collect), or NULL
	cudaStream_t draw_stream = nullptr;
	mm_camera draw_cam;
	const uint32_t *draw_list = nullptr;
	// software rasteriser (mm_raster.cu): texture array, 64-bit key frame buffer, queue of large triangles
	float *d_tex = nullptr;
	uint32_t tex_layers = 0, tex_size = 0;
	unsigned char *d_fb = nullptr;
	size_t fb_pixels = 0;
	unsigned char *d_big = nullptr, *d_big_count = nullptr;
	unsigned long long *h_draw_totals = nullptr, *d_h_draw_totals = nullptr;  // mapped pinned: drawn chunks, vertices i, t

	char err[512] = {0};
};

static inline int mm_fail(mm_ctx *c, int code, const char *fmt, ...) {
	if (c) {
		va_list ap;
		va_start(ap, fmt);
		vsnprintf(c->err, sizeof(c->err), fmt, ap);
		va_end(ap);
	}
	return code;
}

#define MM_CUDA(c, call)                                                                          \
	do {                                                                                          \
		cudaError_t e__ = (call);                                                                 \
		if (e__ != cudaSuccess)                                                                   \
			return mm_fail(c, e__ == cudaErrorMemoryAllocation ? MM_ERR_OOM : MM_ERR_CUDA, "%s: %s", #call, \
			            cudaGetErrorString(e__));                                                 \
	} while (0)

