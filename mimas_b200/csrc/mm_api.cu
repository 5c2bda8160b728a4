// mm_api.cu -- C ABI of libmimas_meshgen_cuda.so (see include/mimas_meshgen_cuda.h).
//
// Host side of the B200 mesher: context, device block table, arenas for the packed
// vertex streams, launch of the analyse + expand kernel pairs, grow-and-rerun when a batch
// does not fit, host<->device staging for the `chunk in, Vec<Vertex> out` crate API.
// There is no CPU implementation of the mesher in this library.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/mimas_meshgen_cuda.h"
#include "mesh_kernel.cuh"
#include "mm_internal.h"
#include "terrain_gen.h"

static_assert(sizeof(mm_vertex) == 36, "Vertex must be 36 bytes (SURVEY.md 8b)");
static_assert(sizeof(mm_range) == 16 && sizeof(mm::Range) == 16, "range layout");
static_assert(sizeof(mm_block_entry) == 12, "block entry layout");
static_assert(mm::kExpandSmemBytes <= 227 * 1024, "the expand CTA must fit the 227 KiB of an SM");

namespace {

#define fail mm_fail

int use_device(mm_ctx *c) {
	MM_CUDA(c, cudaSetDevice(c->device));
	return MM_OK;
}

bool is_device_ptr(const void *p) {
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

bool is_pinned_ptr(const void *p) {
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return a.type == cudaMemoryTypeHost;
}

int ensure_chunk_capacity(mm_ctx *c, mm_slot *sl, uint32_t n) {
	if (n <= sl->cap_chunks) return MM_OK;
	const uint32_t nc = std::max<uint32_t>(n, 1024u);
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	cudaFree(sl->d_ranges_i); cudaFree(sl->d_ranges_t);
	sl->d_ranges_i = sl->d_ranges_t = nullptr;
	sl->cap_chunks = 0;
	MM_CUDA(c, cudaMalloc(&sl->d_ranges_i, sizeof(mm::Range) * (size_t)nc));
	MM_CUDA(c, cudaMalloc(&sl->d_ranges_t, sizeof(mm::Range) * (size_t)nc));
	sl->cap_chunks = nc;
	return MM_OK;
}

int ensure_slot_totals(mm_ctx *c, mm_slot *sl) {  // the two mapped words a slot's totals reach the host through
	if (sl->h_totals) return MM_OK;
	MM_CUDA(c, cudaHostAlloc(&sl->h_totals, sizeof(mm::u64) * 2, cudaHostAllocMapped));
	MM_CUDA(c, cudaHostGetDevicePointer((void **)&sl->d_h_totals, sl->h_totals, 0));
	sl->h_totals[0] = sl->h_totals[1] = 0;
	return MM_OK;
}

void free_slot(mm_slot *sl) {
	cudaFree(sl->d_out_i); cudaFree(sl->d_out_t); cudaFree(sl->d_ranges_i); cudaFree(sl->d_ranges_t);
	if (sl->h_totals) cudaFreeHost(sl->h_totals);
	*sl = mm_slot();
}

int ensure_arena(mm_ctx *c, unsigned char **p, uint64_t *cap, uint64_t quads) {
	if (quads <= *cap) return MM_OK;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	cudaFree(*p);
	*p = nullptr;
	*cap = 0;
	// round up to 1 MiB; +16 B so that a bulk copy never ends past the allocation
	const size_t bytes = (((size_t)quads * mm::kQuadBytes + 16 + (1u << 20) - 1) >> 20) << 20;
	MM_CUDA(c, cudaMalloc(p, bytes));
	*cap = (bytes - 16) / mm::kQuadBytes;
	return MM_OK;
}

int ensure_pinned(mm_ctx *c, unsigned char **p, size_t *have, size_t want) {
	if (want <= *have) return MM_OK;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	if (*p) cudaFreeHost(*p);
	*p = nullptr;
	*have = 0;
	const size_t bytes = ((want + (1u << 20) - 1) >> 20) << 20;
	MM_CUDA(c, cudaHostAlloc(p, bytes, cudaHostAllocDefault));
	*have = bytes;
	return MM_OK;
}

int ensure_in_capacity(mm_ctx *c, uint32_t n) {  // device staging of host inputs
	if (n <= c->cap_in) return MM_OK;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	cudaFree(c->d_in_chunks); cudaFree(c->d_in_offs);
	c->d_in_chunks = nullptr; c->d_in_offs = nullptr; c->cap_in = 0;
	const uint32_t nc = std::max<uint32_t>(n, 256u);
	MM_CUDA(c, cudaMalloc(&c->d_in_chunks, (size_t)nc * 4096));
	MM_CUDA(c, cudaMalloc(&c->d_in_offs, (size_t)nc * 24));
	c->cap_in = nc;
	return MM_OK;
}

// Descriptor slots for launches of up to n chunks (64 KiB per chunk; only the touched lines cost
// bandwidth, and they are re-read from L2 by the expand kernel).
int ensure_desc_capacity(mm_ctx *c, uint32_t n) {
	n = std::min<uint32_t>(n, mm::kMaxLaunchChunks);
	if (n <= c->cap_desc_chunks) return MM_OK;
	const uint32_t nc = std::min<uint32_t>(std::max<uint32_t>(n, 256u), mm::kMaxLaunchChunks);
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	cudaFree(c->d_desc_i); cudaFree(c->d_desc_t);
	c->d_desc_i = c->d_desc_t = nullptr;
	c->cap_desc_chunks = 0;
	MM_CUDA(c, cudaMalloc(&c->d_desc_i, 2 * sizeof(uint32_t) * (size_t)nc * mm::kDescStrideI));
	MM_CUDA(c, cudaMalloc(&c->d_desc_t, 2 * sizeof(uint32_t) * (size_t)nc * mm::kDescStrideT));
	c->cap_desc_chunks = nc;
	return MM_OK;
}

// Enqueue the mesher for (d_chunks, d_offs, n) on the context's stream: per kMaxLaunchChunks
// chunks one analyse + one expand launch, each pair continuing the packed streams where the
// previous one ended.
// Nothing but kernels is enqueued for an untimed batch (the totals reach the host through mapped
// pinned memory), so that back-to-back batches chain by programmatic dependent launch.
struct Targets {  // where a batch's vertices, range records and totals go
	unsigned char *out_i, *out_t;
	uint64_t cap_i, cap_t;
	mm::Range *ranges_i, *ranges_t;
	mm::u64 *host_totals;  // mapped pinned, 2 words
};
Targets slot_targets(const mm_slot *sl) {
	return Targets{sl->d_out_i, sl->d_out_t, sl->cap_i, sl->cap_t, sl->d_ranges_i, sl->d_ranges_t, sl->d_h_totals};
}

// The first kernel of the call carries the completion signal of the newest batch launched before it (see
// AnalyseParams::h_done): once it is past its dependency wait, that batch has landed.
int launch(mm_ctx *c, const unsigned char *d_chunks, const long long *d_offs, const unsigned char *d_halo, uint32_t n,
           bool timed, const Targets &tg, const mm::u64 *carry0 = nullptr, uint32_t *launches = nullptr) {
	int rc;
	if ((rc = ensure_desc_capacity(c, n))) return rc;
	uint32_t launched = 0;
	uint64_t prev_done = c->last_signal_ticket > c->signalled_upto ? c->last_signal_ticket : 0;
	if (timed) MM_CUDA(c, cudaEventRecord(c->ev0, c->stream));
	const mm::u64 *carry = carry0;  // totals of an earlier piece of the same batch, or NULL
	for (uint32_t first = 0; first < n; first += mm::kMaxLaunchChunks) {
		const uint32_t m = std::min<uint32_t>(n - first, mm::kMaxLaunchChunks);
		c->totals_slot ^= 1;
		mm::AnalyseParams a;
		a.chunks = d_chunks + (size_t)first * 4096u;
		a.halo = d_halo ? d_halo + (size_t)first * MM_HALO_BYTES : nullptr;
		a.n = m;
		a.table = c->d_table;
		a.flags = c->d_flags;
		c->pair_parity ^= 1;
		const size_t par = (size_t)c->pair_parity;
		a.desc_i = c->d_desc_i + par * c->cap_desc_chunks * mm::kDescStrideI;
		a.desc_t = c->d_desc_t + par * c->cap_desc_chunks * mm::kDescStrideT;
		a.cnt_i = c->d_cnt + par * 2 * mm::kMaxLaunchChunks;
		a.cnt_t = a.cnt_i + mm::kMaxLaunchChunks;
		a.ticket = c->d_ticket;
		a.ticket_base = c->ticket_base;
		mm::u64 *tr = c->d_trace ? c->d_trace + 8 * (c->trace_next++ % c->trace_pairs) : nullptr;
		a.trace = tr;
		a.h_done = prev_done ? c->d_h_done : nullptr;
		a.prev_done = prev_done;
		if (prev_done) c->signalled_upto = prev_done;
		prev_done = 0;
		cudaLaunchAttribute attr[1];
		attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
		attr[0].val.programmaticStreamSerializationAllowed = c->use_pdl ? 1 : 0;
		// one chunk (the reference's call pattern): one CTA analyses and expands it, one launch
		const bool fused = n == 1 && !d_halo && carry == nullptr && c->fuse_single;
		const uint32_t grid = std::min<uint32_t>(m, (uint32_t)(c->sm_count * c->analyse_ctas_per_sm));
		if (fused) {
			mm::ExpandParams &e = a.x;
			e.offs = d_offs;
			e.n = 1;
			e.table = c->d_table;
			e.desc_i = a.desc_i; e.desc_t = a.desc_t; e.cnt_i = a.cnt_i; e.cnt_t = a.cnt_t;
			e.out_i = tg.out_i; e.out_t = tg.out_t; e.cap_i = tg.cap_i; e.cap_t = tg.cap_t;
			e.ranges_i = tg.ranges_i; e.ranges_t = tg.ranges_t;
			e.carry_in = nullptr;
			e.totals = c->d_ticket + 2 + 2 * c->totals_slot;
			e.host_totals = tg.host_totals;
			e.trace = nullptr;
			e.h_flag = c->flag_wanted ? c->d_h_flag : nullptr;  // mapped path: the caller polls instead of synchronising
			e.flag_value = c->flag_wanted;
			c->flag_armed = c->flag_wanted;
			cudaLaunchConfig_t cfg = {};
			cfg.gridDim = dim3(1);
			cfg.blockDim = dim3(mm::kThreads);
			cfg.stream = c->stream;
			cfg.attrs = attr;
			cfg.numAttrs = 1;
			MM_CUDA(c, (cudaLaunchKernelEx(&cfg, mm::analyse_kernel<false, true>, a)));
			// the fused form draws no tickets
			launched += 1;
			c->total_launches += 1;
			break;
		}
		{
			cudaLaunchConfig_t cfg = {};
			cfg.gridDim = dim3(grid);
			cfg.blockDim = dim3(mm::kThreads);
			cfg.dynamicSmemBytes = 0;
			cfg.stream = c->stream;
			cfg.attrs = attr;
			cfg.numAttrs = 1;
			if (d_halo) MM_CUDA(c, cudaLaunchKernelEx(&cfg, mm::analyse_kernel<true>, a));
			else MM_CUDA(c, cudaLaunchKernelEx(&cfg, mm::analyse_kernel<false>, a));
		}
		c->ticket_base += (mm::u64)m + (mm::u64)grid;  // every CTA draws exactly one ticket >= m

		mm::ExpandParams e;
		e.offs = d_offs + (size_t)first * 3u;
		e.n = m;
		e.table = c->d_table;
		e.desc_i = a.desc_i;
		e.desc_t = a.desc_t;
		e.cnt_i = a.cnt_i;
		e.cnt_t = a.cnt_t;
		e.out_i = tg.out_i;
		e.out_t = tg.out_t;
		e.cap_i = tg.cap_i;
		e.cap_t = tg.cap_t;
		e.ranges_i = tg.ranges_i + first;
		e.ranges_t = tg.ranges_t + first;
		e.carry_in = carry;
		e.totals = c->d_ticket + 2 + 2 * c->totals_slot;
		e.host_totals = tg.host_totals;
		e.trace = tr ? tr + 4 : nullptr;
		e.h_flag = nullptr;
		e.flag_value = 0;
		// programmatic dependent launch (see the kernels): analysis of batch k+1 beside expansion of batch k
		cudaLaunchConfig_t cfg = {};
		cfg.gridDim = dim3((unsigned)c->expand_grid);
		cfg.blockDim = dim3(mm::kExpandThreads);
		cfg.dynamicSmemBytes = mm::kExpandSmemBytes;
		cfg.stream = c->stream;
		cfg.attrs = attr;
		cfg.numAttrs = 1;
		MM_CUDA(c, cudaLaunchKernelEx(&cfg, mm::expand_kernel, e));
		carry = e.totals;
		launched += 2;
		c->total_launches += 2;
	}
	if (timed) MM_CUDA(c, cudaEventRecord(c->ev1, c->stream));
	if (launches) *launches += launched;
	return MM_OK;
}

// Blocks until batch `seq` has landed.  A batch with a later launch behind it is signalled by that launch (mapped host
// word, polled here -- the batches behind it keep running); the newest one is waited for on the stream.
int wait_signal(mm_ctx *c, uint64_t seq) {
	volatile mm::u64 *done = c->h_done;
	if (*done >= seq) return MM_OK;
	if (seq > c->signalled_upto) {  // nothing on the stream carries its signal
		MM_CUDA(c, cudaStreamSynchronize(c->stream));
		return MM_OK;
	}
	for (uint32_t spin = 0;; spin++) {
		if (*done >= seq) return MM_OK;
		if ((spin & 1023u) == 1023u) {
			const cudaError_t e = cudaStreamQuery(c->stream);
			if (e == cudaSuccess) return MM_OK;  // the stream has drained: everything has landed
			if (e != cudaErrorNotReady) return fail(c, MM_ERR_CUDA, "cudaStreamQuery: %s", cudaGetErrorString(e));
		}
		std::this_thread::yield();
	}
}

}  // namespace

extern "C" {

uint32_t mm_abi_version(void) { return MM_ABI_VERSION; }

const char *mm_last_error_string(const mm_ctx *ctx) { return ctx ? ctx->err : "null context"; }

int mm_ctx_create(int device, mm_ctx **out) {
	if (!out) return MM_ERR_INVALID;
	*out = nullptr;
	int count = 0;
	if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
		cudaGetLastError();
		return MM_ERR_NO_DEVICE;  // no CPU fallback by design
	}
	if (device < 0 || device >= count) return MM_ERR_INVALID;
	mm_ctx *c = new (std::nothrow) mm_ctx();
	if (!c) return MM_ERR_OOM;
	c->device = device;
	int rc = MM_OK;
	auto body = [&]() -> int {
		MM_CUDA(c, cudaSetDevice(device));
		cudaDeviceProp prop;
		MM_CUDA(c, cudaGetDeviceProperties(&prop, device));
		if (prop.major < 10) return fail(c, MM_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a", device, prop.major, prop.minor);
		c->sm_count = prop.multiProcessorCount;
		MM_CUDA(c, cudaFuncSetAttribute(mm::analyse_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
		MM_CUDA(c, cudaFuncSetAttribute(mm::analyse_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
		MM_CUDA(c, cudaFuncSetAttribute(mm::expand_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mm::kExpandSmemBytes));
		MM_CUDA(c, cudaFuncSetAttribute(mm::expand_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
		int occ = 0;
		MM_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, mm::analyse_kernel<false>, mm::kThreads, 0));
		int occh = 0;  // the halo instantiation runs on the same grid
		MM_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occh, mm::analyse_kernel<true>, mm::kThreads, 0));
		occ = std::min(occ, occh);
		if (occ < 1) return fail(c, MM_ERR_CUDA, "analyse kernel does not fit on an SM");
		c->ctas_per_sm = occ;
		c->analyse_ctas_per_sm = occ;
		if (const char *e = getenv("MM_ANALYSE_GRID_PER_SM")) c->analyse_ctas_per_sm = std::max(1, std::min(occ, atoi(e)));
		int occ2 = 0;
		MM_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, mm::expand_kernel, mm::kExpandThreads, mm::kExpandSmemBytes));
		if (occ2 < 1) return fail(c, MM_ERR_CUDA, "expand kernel does not fit on an SM");
		c->expand_grid = c->sm_count;  // one expanding CTA per SM, whatever would fit: the rest of the SM is the analysis's
		c->use_pdl = getenv("MM_NO_PDL") == nullptr;
		c->fuse_single = getenv("MM_NO_FUSED_SINGLE") == nullptr;
		if (const char *e = getenv("MM_PIPE_MIN_CHUNKS")) c->pipe_min_chunks = (uint32_t)std::max(0, atoi(e));
		c->no_small = getenv("MM_NO_SMALL_PATH") != nullptr;
		if (const char *e = getenv("MM_MAPPED_MAX_CHUNKS")) c->mapped_max_chunks = (uint32_t)strtoul(e, nullptr, 10);
		MM_CUDA(c, cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
		c->stream = c->own_stream;
		MM_CUDA(c, cudaEventCreate(&c->ev0));
		MM_CUDA(c, cudaEventCreate(&c->ev1));
		MM_CUDA(c, cudaMalloc(&c->d_table, sizeof(mm::TabEntry) * 256));
		MM_CUDA(c, cudaMalloc(&c->d_ticket, sizeof(mm::u64) * 8));
		MM_CUDA(c, cudaMemset(c->d_ticket, 0, sizeof(mm::u64) * 8));
		MM_CUDA(c, cudaMalloc(&c->d_flags, 256));
		MM_CUDA(c, cudaMalloc(&c->d_cnt, sizeof(uint32_t) * 4 * mm::kMaxLaunchChunks));
		MM_CUDA(c, cudaMemset(c->d_cnt, 0, sizeof(uint32_t) * 4 * mm::kMaxLaunchChunks));
		{
			const int rs = ensure_slot_totals(c, &c->ring[0]);
			if (rs) return rs;
		}
		// the host paths (mapped / pipelined) read their totals from words of their own: a ring batch still in flight keeps its slot's
		MM_CUDA(c, cudaHostAlloc(&c->h_host_totals, sizeof(mm::u64) * 2, cudaHostAllocMapped));
		MM_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_h_host_totals, c->h_host_totals, 0));
		MM_CUDA(c, cudaHostAlloc(&c->h_flag, sizeof(mm::u64), cudaHostAllocMapped));
		MM_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_h_flag, c->h_flag, 0));
		*c->h_flag = 0;
		MM_CUDA(c, cudaHostAlloc(&c->h_done, sizeof(mm::u64), cudaHostAllocMapped));
		MM_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_h_done, c->h_done, 0));
		*c->h_done = 0;
		{  // consumers get the SMs first whenever the persistent mesher kernels hand some back
			int lo = 0, hi = 0;
			MM_CUDA(c, cudaDeviceGetStreamPriorityRange(&lo, &hi));
			MM_CUDA(c, cudaStreamCreateWithPriority(&c->s_draw, cudaStreamNonBlocking, hi));
		}
		return MM_OK;
	};
	rc = body();
	if (rc != MM_OK) {
		fprintf(stderr, "mm_ctx_create: %s\n", c->err);
		mm_ctx_destroy(c);
		return rc;
	}
	*out = c;
	return MM_OK;
}

int mm_ctx_destroy(mm_ctx *c) {
	if (!c) return MM_ERR_INVALID;
	cudaSetDevice(c->device);
	if (c->own_stream) cudaStreamSynchronize(c->own_stream);
	if (c->s_draw) cudaStreamSynchronize(c->s_draw);
	for (uint32_t k = 0; k < kMaxRingSlots; k++) free_slot(&c->ring[k]);
	cudaFree(c->d_table); cudaFree(c->d_cnt); cudaFree(c->d_desc_i); cudaFree(c->d_desc_t); cudaFree(c->d_flags); cudaFree(c->d_ticket);
	if (c->h_done) cudaFreeHost(c->h_done);
	if (c->h_host_totals) cudaFreeHost(c->h_host_totals);
	if (c->h_flag) cudaFreeHost(c->h_flag);
	if (c->s_draw) cudaStreamDestroy(c->s_draw);
	cudaFree(c->d_in_chunks); cudaFree(c->d_in_offs); cudaFree(c->d_in_halo); cudaFree(c->d_trace);
	if (c->h_in) cudaFreeHost(c->h_in);
	if (c->h_small) cudaFreeHost(c->h_small);
	if (c->h_draw_totals) cudaFreeHost(c->h_draw_totals);
	cudaFree(c->d_cull); cudaFree(c->d_shaded);
	cudaFree(c->d_tex); cudaFree(c->d_fb); cudaFree(c->d_big); cudaFree(c->d_big_count);
	if (c->ev2) cudaEventDestroy(c->ev2);
	if (c->ev3) cudaEventDestroy(c->ev3);
	if (c->h_out) cudaFreeHost(c->h_out);
	if (c->s_in) cudaStreamDestroy(c->s_in);
	if (c->s_out) cudaStreamDestroy(c->s_out);
	for (cudaEvent_t e : c->ev_in)
		if (e) cudaEventDestroy(e);
	if (c->ev0) cudaEventDestroy(c->ev0);
	if (c->ev1) cudaEventDestroy(c->ev1);
	if (c->own_stream) cudaStreamDestroy(c->own_stream);
	delete c;
	return MM_OK;
}

int mm_ctx_set_stream(mm_ctx *c, void *cuda_stream) {
	if (!c) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
	return MM_OK;
}

int mm_ctx_sync(mm_ctx *c) {
	if (!c) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	return MM_OK;
}

int mm_device_info(mm_ctx *c, int *sm_count, int *ctas_per_sm, int *threads_per_cta) {
	if (!c) return MM_ERR_INVALID;
	if (sm_count) *sm_count = c->sm_count;
	if (ctas_per_sm) *ctas_per_sm = c->ctas_per_sm;
	if (threads_per_cta) *threads_per_cta = mm::kThreads;
	return MM_OK;
}

int mm_launch_count(mm_ctx *c, uint64_t *launches) {
	if (!c || !launches) return MM_ERR_INVALID;
	*launches = c->total_launches;
	return MM_OK;
}

int mm_host_alloc(mm_ctx *c, size_t bytes, void **h_ptr) {
	if (!c || !h_ptr) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	*h_ptr = nullptr;
	MM_CUDA(c, cudaHostAlloc(h_ptr, bytes ? bytes : 1, cudaHostAllocDefault));
	return MM_OK;
}
int mm_host_free(mm_ctx *c, void *h_ptr) {
	if (!c) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	MM_CUDA(c, cudaFreeHost(h_ptr));
	return MM_OK;
}

int mm_table_from_cache(const uint16_t fb[3], const mm_block_entry *known, uint32_t n_known,
                        mm_block_entry out[MM_TABLE_ENTRIES]) {
	if (!fb || !out || (n_known && !known) || n_known > MM_TABLE_ENTRIES) return MM_ERR_INVALID;
	for (uint32_t i = 0; i < MM_TABLE_ENTRIES; i++) {
		mm_block_entry e;
		memset(&e, 0, sizeof(e));
		if (i < n_known) {
			e = known[i];
			if (e.kind > MM_KIND_CROSSED) return MM_ERR_INVALID;
		} else {  // get_bl_tex_ids: unknown id -> Some(fallback) (lib.rs:105-109); never crossed (lib.rs:110-113)
			e.kind = MM_KIND_BLOCKY;
			e.id_sides = fb[0]; e.id_top = fb[1]; e.id_bottom = fb[2];
		}
		out[i] = e;
	}
	return MM_OK;
}

int mm_set_block_table(mm_ctx *c, const mm_block_entry table[MM_TABLE_ENTRIES]) {
	if (!c || !table) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	mm::TabEntry t[256];
	unsigned char fl[256];
	for (int i = 0; i < 256; i++) {
		const mm_block_entry &e = table[i];
		if (e.kind > MM_KIND_CROSSED) return fail(c, MM_ERR_INVALID, "table[%d].kind = %u", i, e.kind);
		if (e.kind == MM_KIND_CROSSED) t[i].x = e.id_crossed;
		else t[i].x = (uint32_t)e.id_sides | ((uint32_t)e.id_top << 16);
		t[i].y = (uint32_t)e.id_bottom | ((uint32_t)e.kind << 16);
		if (e.kind == MM_KIND_NONE) { t[i].x = 0; t[i].y = 0; }
	}
	mm::build_flags(t, fl);
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	MM_CUDA(c, cudaMemcpy(c->d_table, t, sizeof(t), cudaMemcpyHostToDevice));
	MM_CUDA(c, cudaMemcpy(c->d_flags, fl, sizeof(fl), cudaMemcpyHostToDevice));
	c->have_table = true;
	return MM_OK;
}

int mm_reserve(mm_ctx *c, uint32_t max_chunks, uint64_t iv, uint64_t tv) {
	if (!c) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	int rc;
	for (uint32_t k = 0; k < c->ring_slots; k++) {  // every slot of the ring: a batch may land in any of them
		mm_slot *sl = &c->ring[k];
		if ((rc = ensure_chunk_capacity(c, sl, max_chunks))) return rc;
		if ((rc = ensure_arena(c, &sl->d_out_i, &sl->cap_i, (iv + 5) / 6))) return rc;
		if ((rc = ensure_arena(c, &sl->d_out_t, &sl->cap_t, (tv + 5) / 6))) return rc;
	}
	return MM_OK;
}

int mm_ring_configure(mm_ctx *c, uint32_t slots) {
	if (!c || slots < 1 || slots > kMaxRingSlots) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	int rc;
	for (uint32_t k = slots; k < kMaxRingSlots; k++) free_slot(&c->ring[k]);
	for (uint32_t k = 0; k < slots; k++) {
		mm_slot *sl = &c->ring[k];
		if ((rc = ensure_slot_totals(c, sl))) return rc;
		sl->ticket = 0;  // outstanding tickets do not survive a reconfiguration
		sl->pending = sl->finished = false;
	}
	c->ring_slots = slots;
	c->cur = &c->ring[0];
	return MM_OK;
}

// One batch into the next slot of the ring; returns its ticket.
static int submit(mm_ctx *c, const uint8_t *d_chunks, const int64_t *d_offs, const uint8_t *d_halo, uint32_t n, uint64_t *ticket) {
	if (!c || (n && (!d_chunks || !d_offs))) return MM_ERR_INVALID;
	// the chunks are fetched by bulk async copies, the halo rows as 16-byte loads
	if (n && ((uintptr_t)d_chunks & 15u)) return fail(c, MM_ERR_INVALID, "chunks must be 16-byte aligned");
	if (n && ((uintptr_t)d_offs & 7u)) return fail(c, MM_ERR_INVALID, "offs must be 8-byte aligned");
	if (d_halo && ((uintptr_t)d_halo & 15u)) return fail(c, MM_ERR_INVALID, "halo must be 16-byte aligned");
	if (!c->have_table) return fail(c, MM_ERR_NO_TABLE, "mm_set_block_table has not been called");
	if (use_device(c)) return MM_ERR_CUDA;
	const uint64_t t = c->next_ticket;
	mm_slot *sl = &c->ring[t % c->ring_slots];
	int rc;
	if ((rc = ensure_slot_totals(c, sl))) return rc;
	if (n && (rc = ensure_chunk_capacity(c, sl, n))) return rc;
	sl->finished = false;
	sl->pending = false;
	sl->last_chunks = d_chunks;
	sl->last_offs = (const long long *)d_offs;
	sl->last_halo = d_halo;
	sl->last_n = n;
	sl->last_launches = 0;
	sl->tot_i = sl->tot_t = 0;
	sl->last_timed = c->time_next;
	sl->wait_ticket = c->last_signal_ticket;  // an empty batch is complete when everything before it is
	if (n) {
		if ((rc = launch(c, d_chunks, sl->last_offs, d_halo, n, c->time_next, slot_targets(sl), nullptr, &sl->last_launches))) return rc;
		sl->wait_ticket = ++c->last_signal_ticket;  // completion signals count on their own: a re-run takes one too
	}
	c->time_next = false;
	sl->ticket = t;
	sl->pending = true;
	c->cur = sl;
	c->next_ticket = t + 1;
	if (ticket) *ticket = t;
	return MM_OK;
}

int mm_mesh_batch_async(mm_ctx *c, const uint8_t *d_chunks, const int64_t *d_offs, uint32_t n) {
	return submit(c, d_chunks, d_offs, nullptr, n, nullptr);
}

int mm_mesh_batch_halo_async(mm_ctx *c, const uint8_t *d_chunks, const int64_t *d_offs, const uint8_t *d_halo, uint32_t n) {
	return submit(c, d_chunks, d_offs, d_halo, n, nullptr);
}

int mm_mesh_batch_submit(mm_ctx *c, const uint8_t *d_chunks, const int64_t *d_offs, const uint8_t *d_halo, uint32_t n,
                         uint64_t *ticket) {
	if (!ticket) return MM_ERR_INVALID;
	return submit(c, d_chunks, d_offs, d_halo, n, ticket);
}

int mm_mesh_batches_submit(mm_ctx *c, const uint8_t *const *d_chunks, const int64_t *const *d_offs, const uint32_t *n,
                           uint32_t k, uint32_t repeat, uint64_t *first_ticket) {
	if (!c || !d_chunks || !d_offs || !n || !k) return MM_ERR_INVALID;
	uint64_t first = 0;
	for (uint32_t r = 0; r < std::max<uint32_t>(repeat, 1u); r++)
		for (uint32_t i = 0; i < k; i++) {
			uint64_t t = 0;
			const int rc = submit(c, d_chunks[i], d_offs[i], nullptr, n[i], &t);
			if (rc) return rc;
			if (!first) first = t;
		}
	if (first_ticket) *first_ticket = first;
	return MM_OK;
}

static void fill_batch_out(const mm_ctx *, const mm_slot *sl, mm_batch_out *out) {
	memset(out, 0, sizeof(*out));
	out->d_intransparent = (const mm_vertex *)sl->d_out_i;
	out->d_transparent = (const mm_vertex *)sl->d_out_t;
	out->d_ranges_i = (const mm_range *)sl->d_ranges_i;
	out->d_ranges_t = (const mm_range *)sl->d_ranges_t;
	out->n_intransparent = sl->tot_i * 6;
	out->n_transparent = sl->tot_t * 6;
	out->n_chunks = sl->last_n;
	out->launches = sl->last_launches;
	out->kernel_ms = sl->last_ms;
}

// Completes the batch a slot holds: waits for its kernels, and if the slot's arenas were too small grows them
// and runs the batch again (its inputs must still be resident).
static int finish_slot(mm_ctx *c, mm_slot *sl) {
	if (sl->last_n) {
		for (int attempt = 0;; attempt++) {
			int rc;
			if ((rc = wait_signal(c, sl->wait_ticket))) return rc;
			sl->tot_i = sl->h_totals[0];
			sl->tot_t = sl->h_totals[1];
			if (sl->tot_i <= sl->cap_i && sl->tot_t <= sl->cap_t) break;
			if (attempt >= 1) return fail(c, MM_ERR_CUDA, "arena still too small after growth");
			// grow to fit (+1/8 head room) and run again: the inputs are still resident
			if ((rc = ensure_arena(c, &sl->d_out_i, &sl->cap_i, sl->tot_i + sl->tot_i / 8))) return rc;
			if ((rc = ensure_arena(c, &sl->d_out_t, &sl->cap_t, sl->tot_t + sl->tot_t / 8))) return rc;
			if ((rc = launch(c, sl->last_chunks, sl->last_offs, sl->last_halo, sl->last_n, sl->last_timed, slot_targets(sl), nullptr,
			                 &sl->last_launches)))
				return rc;
			sl->wait_ticket = ++c->last_signal_ticket;
		}
		sl->last_ms = 0.f;  // an untimed (mm_mesh_batch_async) batch enqueues kernels only
		if (sl->last_timed) {
			MM_CUDA(c, cudaEventSynchronize(c->ev1));
			MM_CUDA(c, cudaEventElapsedTime(&sl->last_ms, c->ev0, c->ev1));
		}
	} else {
		int rc;
		if ((rc = wait_signal(c, sl->wait_ticket))) return rc;
		sl->last_ms = 0.f;
	}
	sl->pending = false;
	sl->finished = true;
	return MM_OK;
}

int mm_batch_finish(mm_ctx *c, mm_batch_out *out) {
	if (!c) return MM_ERR_INVALID;
	mm_slot *sl = c->cur;
	if (!sl->pending) return fail(c, MM_ERR_NO_BATCH, "no batch in flight");
	if (use_device(c)) return MM_ERR_CUDA;
	int rc;
	if ((rc = finish_slot(c, sl))) return rc;
	if (out) fill_batch_out(c, sl, out);
	return MM_OK;
}

int mm_batch_wait(mm_ctx *c, uint64_t ticket, mm_batch_out *out) {
	if (!c || !ticket) return MM_ERR_INVALID;
	mm_slot *sl = &c->ring[ticket % c->ring_slots];
	if (sl->ticket != ticket) return fail(c, MM_ERR_NO_BATCH, "ticket %llu is not in the ring (slot reused or never submitted)", (unsigned long long)ticket);
	if (use_device(c)) return MM_ERR_CUDA;
	int rc;
	if (sl->pending && (rc = finish_slot(c, sl))) return rc;
	if (out) fill_batch_out(c, sl, out);
	return MM_OK;
}

int mm_batch_poll(mm_ctx *c, uint64_t ticket, int *done) {
	if (!c || !ticket || !done) return MM_ERR_INVALID;
	const mm_slot *sl = &c->ring[ticket % c->ring_slots];
	if (sl->ticket != ticket) return fail(c, MM_ERR_NO_BATCH, "ticket %llu is not in the ring", (unsigned long long)ticket);
	*done = !sl->pending || *(volatile mm::u64 *)c->h_done >= sl->wait_ticket ||
	        (sl->wait_ticket > c->signalled_upto && cudaStreamQuery(c->stream) == cudaSuccess);
	return MM_OK;
}

int mm_mesh_batch(mm_ctx *c, const uint8_t *chunks, const int64_t *offs, const uint8_t *halo, uint32_t n,
                  mm_batch_out *out) {
	if (!c || (n && (!chunks || !offs))) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	const uint8_t *dc = chunks;
	const int64_t *dof = offs;
	const uint8_t *dh = halo;
	if (n && halo && !is_device_ptr(halo)) {
		if (n > c->cap_in_halo) {
			MM_CUDA(c, cudaStreamSynchronize(c->stream));
			cudaFree(c->d_in_halo);
			c->d_in_halo = nullptr; c->cap_in_halo = 0;
			const uint32_t nc = std::max<uint32_t>(n, 256u);
			MM_CUDA(c, cudaMalloc(&c->d_in_halo, (size_t)nc * MM_HALO_BYTES));
			c->cap_in_halo = nc;
		}
		MM_CUDA(c, cudaMemcpyAsync(c->d_in_halo, halo, (size_t)n * MM_HALO_BYTES, cudaMemcpyHostToDevice, c->stream));
		dh = c->d_in_halo;
	}
	if (n && (!is_device_ptr(chunks) || !is_device_ptr(offs))) {
		int rc0;
		if ((rc0 = ensure_in_capacity(c, n))) return rc0;
		if (!is_device_ptr(chunks)) {
			MM_CUDA(c, cudaMemcpyAsync(c->d_in_chunks, chunks, (size_t)n * 4096, cudaMemcpyHostToDevice, c->stream));
			dc = c->d_in_chunks;
		}
		if (!is_device_ptr(offs)) {
			MM_CUDA(c, cudaMemcpyAsync(c->d_in_offs, offs, (size_t)n * 24, cudaMemcpyHostToDevice, c->stream));
			dof = (const int64_t *)c->d_in_offs;
		}
	}
	int rc;
	c->time_next = true;  // the synchronous call brackets its kernels with events: mm_batch_out.kernel_ms
	if ((rc = submit(c, dc, dof, dh, n, nullptr))) {
		c->time_next = false;
		return rc;
	}
	return mm_batch_finish(c, out);
}

// Copies the batch of slot `sl` to the host (any pointer may be NULL) on `stream` and waits for the copies.
static int copy_slot(mm_ctx *c, const mm_slot *sl, cudaStream_t stream, mm_vertex *vi, mm_vertex *vt, mm_range *ri, mm_range *rt) {
	if (vi && sl->tot_i) MM_CUDA(c, cudaMemcpyAsync(vi, sl->d_out_i, sl->tot_i * mm::kQuadBytes, cudaMemcpyDeviceToHost, stream));
	if (vt && sl->tot_t) MM_CUDA(c, cudaMemcpyAsync(vt, sl->d_out_t, sl->tot_t * mm::kQuadBytes, cudaMemcpyDeviceToHost, stream));
	if (ri && sl->last_n) MM_CUDA(c, cudaMemcpyAsync(ri, sl->d_ranges_i, (size_t)sl->last_n * 16, cudaMemcpyDeviceToHost, stream));
	if (rt && sl->last_n) MM_CUDA(c, cudaMemcpyAsync(rt, sl->d_ranges_t, (size_t)sl->last_n * 16, cudaMemcpyDeviceToHost, stream));
	MM_CUDA(c, cudaStreamSynchronize(stream));
	return MM_OK;
}

int mm_copy_out(mm_ctx *c, mm_vertex *vi, mm_vertex *vt, mm_range *ri, mm_range *rt) {
	if (!c) return MM_ERR_INVALID;
	if (!c->cur->finished) return fail(c, MM_ERR_NO_BATCH, "no finished batch");
	if (use_device(c)) return MM_ERR_CUDA;
	return copy_slot(c, c->cur, c->stream, vi, vt, ri, rt);
}

int mm_copy_out_ticket(mm_ctx *c, uint64_t ticket, mm_vertex *vi, mm_vertex *vt, mm_range *ri, mm_range *rt) {
	if (!c || !ticket) return MM_ERR_INVALID;
	mm_slot *sl = &c->ring[ticket % c->ring_slots];
	if (sl->ticket != ticket) return fail(c, MM_ERR_NO_BATCH, "ticket %llu is not in the ring", (unsigned long long)ticket);
	if (use_device(c)) return MM_ERR_CUDA;
	int rc;
	if (sl->pending && (rc = finish_slot(c, sl))) return rc;
	// on the consumers' stream: the batches queued behind this one keep meshing meanwhile
	return copy_slot(c, sl, c->s_draw, vi, vt, ri, rt);
}

// The mapped path.  Small batches (the reference's own call pattern is one chunk per call,
// client.rs:259-267) skip every cudaMemcpy: the kernels read the chunks from, and write vertices and
// range records to, mapped pinned host memory.  One memcpy in, two launches, one stream synchronise.
// The mapped block grows on demand (arenas that turn out too small: grow and run again, like the device
// arenas) up to kMappedMaxBytes; beyond that, or above c->mapped_max_chunks chunks, *done = false and the
// caller takes the copy path.
constexpr size_t kMappedMaxBytes = (size_t)1 << 30;
struct MappedLayout {
	size_t chunks, offs, ranges_i, ranges_t, out_i, out_t, total;
	MappedLayout(uint32_t nc, uint64_t qi, uint64_t qt) {
		const auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
		chunks = 0;
		offs = up(chunks + (size_t)nc * 4096);
		ranges_i = up(offs + (size_t)nc * 24);
		ranges_t = up(ranges_i + (size_t)nc * 16);
		out_i = up(ranges_t + (size_t)nc * 16);
		out_t = up(out_i + qi * mm::kQuadBytes + 16);
		total = out_t + qt * mm::kQuadBytes + 16;
	}
};
static int small_batch_host(mm_ctx *c, const uint8_t *chunks, const int64_t *offs, uint32_t n, mm_host_mesh *out, bool *done) {
	*done = false;
	if (!c->have_table) return fail(c, MM_ERR_NO_TABLE, "mm_set_block_table has not been called");
	int rc;
	bool regrow = false;
	for (int attempt = 0;; attempt++) {
		if (!c->h_small || n > c->mapped_chunks || regrow) {  // first use, more chunks than ever before, or bigger arenas
			const uint32_t nc = std::max<uint32_t>(std::max<uint32_t>(n, 32u), c->mapped_chunks);
			const uint64_t qi = std::max<uint64_t>(c->mapped_qi, 49152), qt = std::max<uint64_t>(c->mapped_qt, 16384);  // 10.1 + 3.4 MiB
			const MappedLayout L(nc, qi, qt);
			if (L.total > kMappedMaxBytes) return MM_OK;
			MM_CUDA(c, cudaStreamSynchronize(c->stream));
			if (c->h_small) cudaFreeHost(c->h_small);
			c->h_small = c->d_small = nullptr;
			MM_CUDA(c, cudaHostAlloc(&c->h_small, L.total, cudaHostAllocMapped));
			MM_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_small, c->h_small, 0));
			c->mapped_chunks = nc; c->mapped_qi = qi; c->mapped_qt = qt;
			regrow = false;
		}
		const MappedLayout L(c->mapped_chunks, c->mapped_qi, c->mapped_qt);
		// inputs in page-locked memory (mm_host_alloc) are read in place; pageable ones are copied into the block
		const unsigned char *d_chunks = nullptr;
		const long long *d_offs = nullptr;
		if (((uintptr_t)chunks & 15u) == 0 && ((uintptr_t)offs & 7u) == 0 && is_pinned_ptr(chunks) && is_pinned_ptr(offs) &&
		    cudaHostGetDevicePointer((void **)&d_chunks, (void *)chunks, 0) == cudaSuccess &&
		    cudaHostGetDevicePointer((void **)&d_offs, (void *)offs, 0) == cudaSuccess) {
		} else {
			cudaGetLastError();
			memcpy(c->h_small + L.chunks, chunks, (size_t)n * 4096);
			memcpy(c->h_small + L.offs, offs, (size_t)n * 24);
			d_chunks = c->d_small + L.chunks;
			d_offs = (const long long *)(c->d_small + L.offs);
		}
		// nothing of this call touches a ring slot: vertices, range records and totals all go to mapped host memory
		Targets tg{c->d_small + L.out_i, c->d_small + L.out_t, c->mapped_qi, c->mapped_qt,
		           (mm::Range *)(c->d_small + L.ranges_i), (mm::Range *)(c->d_small + L.ranges_t), c->d_h_host_totals};
		// one chunk = one fused launch: it tells a mapped host word when its vertices have landed, and polling that
		// word costs a few microseconds less than synchronising the stream
		c->flag_wanted = ++c->flag_seq;
		c->flag_armed = 0;
		rc = launch(c, d_chunks, d_offs, nullptr, n, false, tg);
		c->flag_wanted = 0;
		if (rc) return rc;
		bool landed = false;
		if (c->flag_armed) {
			const auto t0 = std::chrono::steady_clock::now();
			for (uint32_t spin = 0;; spin++) {
				if (*(volatile mm::u64 *)c->h_flag >= c->flag_armed) { landed = true; break; }
				if ((spin & 255u) == 255u && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(2)) break;
#if defined(__x86_64__)
				__builtin_ia32_pause();
#endif
			}
		}
		if (!landed) MM_CUDA(c, cudaStreamSynchronize(c->stream));
		const uint64_t ti = c->h_host_totals[0], tt = c->h_host_totals[1];
		if (ti > c->mapped_qi || tt > c->mapped_qt) {  // nothing was written: grow the mapped arenas and run again
			if (attempt >= 1) return MM_OK;
			const MappedLayout G(c->mapped_chunks, std::max(c->mapped_qi, ti + ti / 8), std::max(c->mapped_qt, tt + tt / 8));
			if (G.total > kMappedMaxBytes) return MM_OK;
			c->mapped_qi = std::max(c->mapped_qi, ti + ti / 8);
			c->mapped_qt = std::max(c->mapped_qt, tt + tt / 8);
			regrow = true;
			continue;
		}
		out->intransparent = (const mm_vertex *)(c->h_small + L.out_i);
		out->transparent = (const mm_vertex *)(c->h_small + L.out_t);
		out->ranges_i = (const mm_range *)(c->h_small + L.ranges_i);
		out->ranges_t = (const mm_range *)(c->h_small + L.ranges_t);
		out->n_intransparent = ti * 6;
		out->n_transparent = tt * 6;
		out->n_chunks = n;
		out->path = MM_PATH_MAPPED;
		*done = true;
		return MM_OK;
	}
}

// The pipelined path.  A large host batch is PCIe-bound (per chunk 4 KiB go in, some 43 KiB of vertices
// come back), so the batch is cut into pieces: all pieces are sent on one stream, each is meshed as soon as
// it has arrived -- continuing the packed streams of the piece before it -- and its share of the vertex
// streams starts its way back on a third stream while the next piece is meshed.  Both DMA directions and
// the kernels overlap; the call costs little more than the device-to-host copy alone.  Inputs must be
// page-locked (the caller stages pageable ones).  *done = false: the arenas are not sized yet or turned out
// too small; the caller takes the plain path, which grows them.
static int pipelined_batch_host(mm_ctx *c, const uint8_t *hc, const int64_t *ho, uint32_t n, mm_host_mesh *out, bool *done) {
	*done = false;
	if (!c->have_table) return fail(c, MM_ERR_NO_TABLE, "mm_set_block_table has not been called");
	mm_slot *sl = c->cur;
	if (!sl->cap_i && !sl->cap_t) return MM_OK;
	if (c->host_qi_per_chunk < 0.0) return MM_OK;  // no earlier host batch to size the pinned result by
	int rc;
	if ((rc = ensure_in_capacity(c, n))) return rc;
	if ((rc = ensure_chunk_capacity(c, sl, n))) return rc;
	const auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
	const size_t br = (size_t)n * 16;
	// room in the pinned result: what the last host batch needed per chunk plus a quarter, never more than the
	// device arenas hold (those may be far larger after a big device-resident batch)
	const uint64_t room_i = std::min<uint64_t>(sl->cap_i, (uint64_t)(c->host_qi_per_chunk * 1.25 * n) + 4096);
	const uint64_t room_t = std::min<uint64_t>(sl->cap_t, (uint64_t)(c->host_qt_per_chunk * 1.25 * n) + 4096);
	const size_t oi = 0, ot = up(room_i * mm::kQuadBytes), ori = up(ot + room_t * mm::kQuadBytes), ort = ori + up(br);
	if ((rc = ensure_pinned(c, &c->h_out, &c->h_out_bytes, ort + br + 256))) return rc;
	if (!c->s_in) {
		MM_CUDA(c, cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
		MM_CUDA(c, cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
		for (cudaEvent_t &e : c->ev_in) MM_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	}
	MM_CUDA(c, cudaStreamSynchronize(c->stream));  // earlier batches are done with the staging buffers and the arenas
	const uint32_t piece = std::min<uint32_t>(mm::kMaxLaunchChunks, std::max<uint32_t>(512u, ((n + 7u) / 8u + 255u) & ~255u));
	const uint32_t pieces = (n + piece - 1) / piece;
	const uint32_t ahead = 8;  // pieces sent before their kernels are enqueued (one event each)
	auto send = [&](uint32_t k) -> int {
		const uint32_t first = k * piece, m = std::min(piece, n - first);
		MM_CUDA(c, cudaMemcpyAsync(c->d_in_chunks + (size_t)first * 4096, hc + (size_t)first * 4096, (size_t)m * 4096, cudaMemcpyHostToDevice, c->s_in));
		MM_CUDA(c, cudaMemcpyAsync(c->d_in_offs + (size_t)first * 3, ho + (size_t)first * 3, (size_t)m * 24, cudaMemcpyHostToDevice, c->s_in));
		MM_CUDA(c, cudaEventRecord(c->ev_in[k % ahead], c->s_in));
		return MM_OK;
	};
	for (uint32_t k = 0; k < std::min(pieces, ahead); k++)
		if ((rc = send(k))) return rc;
	sl->pending = sl->finished = false;
	sl->ticket = 0;
	sl->last_launches = 0;
	const mm::u64 *carry = nullptr;
	uint64_t pi = 0, pt = 0;  // quads already on their way back
	bool fits = true;
	for (uint32_t k = 0; k < pieces && fits; k++) {
		const uint32_t first = k * piece, m = std::min(piece, n - first);
		MM_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_in[k % ahead], 0));
		const Targets tg{sl->d_out_i, sl->d_out_t, sl->cap_i, sl->cap_t, sl->d_ranges_i + first, sl->d_ranges_t + first, c->d_h_host_totals};
		if ((rc = launch(c, c->d_in_chunks + (size_t)first * 4096, c->d_in_offs + (size_t)first * 3, nullptr, m, false, tg, carry,
		                 &sl->last_launches)))
			return rc;
		carry = c->d_ticket + 2 + 2 * c->totals_slot;
		MM_CUDA(c, cudaStreamSynchronize(c->stream));
		if (k + ahead < pieces && (rc = send(k + ahead))) return rc;  // its event is free again
		const uint64_t ti = c->h_host_totals[0], tt = c->h_host_totals[1];
		if (ti > room_i || tt > room_t) {
			fits = false;
			break;
		}
		if (ti > pi) MM_CUDA(c, cudaMemcpyAsync(c->h_out + oi + pi * mm::kQuadBytes, sl->d_out_i + pi * mm::kQuadBytes, (ti - pi) * mm::kQuadBytes, cudaMemcpyDeviceToHost, c->s_out));
		if (tt > pt) MM_CUDA(c, cudaMemcpyAsync(c->h_out + ot + pt * mm::kQuadBytes, sl->d_out_t + pt * mm::kQuadBytes, (tt - pt) * mm::kQuadBytes, cudaMemcpyDeviceToHost, c->s_out));
		pi = ti;
		pt = tt;
	}
	if (!fits) {  // the plain path grows the arenas and runs the whole batch again
		MM_CUDA(c, cudaStreamSynchronize(c->s_in));
		MM_CUDA(c, cudaStreamSynchronize(c->s_out));
		return MM_OK;
	}
	MM_CUDA(c, cudaMemcpyAsync(c->h_out + ori, sl->d_ranges_i, br, cudaMemcpyDeviceToHost, c->s_out));
	MM_CUDA(c, cudaMemcpyAsync(c->h_out + ort, sl->d_ranges_t, br, cudaMemcpyDeviceToHost, c->s_out));
	MM_CUDA(c, cudaStreamSynchronize(c->s_out));
	sl->tot_i = pi;
	sl->tot_t = pt;
	sl->last_chunks = c->d_in_chunks;
	sl->last_offs = c->d_in_offs;
	sl->last_halo = nullptr;
	sl->last_n = n;
	sl->last_ms = 0.f;
	sl->last_timed = false;
	sl->finished = true;  // the device arenas hold this batch: mm_copy_out works
	c->host_qi_per_chunk = (double)pi / n;
	c->host_qt_per_chunk = (double)pt / n;
	out->intransparent = (const mm_vertex *)(c->h_out + oi);
	out->transparent = (const mm_vertex *)(c->h_out + ot);
	out->ranges_i = (const mm_range *)(c->h_out + ori);
	out->ranges_t = (const mm_range *)(c->h_out + ort);
	out->n_intransparent = pi * 6;
	out->n_transparent = pt * 6;
	out->n_chunks = n;
	out->path = MM_PATH_PIPELINED;
	*done = true;
	return MM_OK;
}

int mm_mesh_batch_host(mm_ctx *c, const uint8_t *chunks, const int64_t *offs, uint32_t n, mm_host_mesh *out) {
	if (!c || !out || (n && (!chunks || !offs))) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	int rc;
	if (n && n <= c->mapped_max_chunks && !c->no_small) {
		bool done = false;
		if ((rc = small_batch_host(c, chunks, offs, n, out, &done))) return rc;
		if (done) return MM_OK;
	}
	// pageable inputs are staged through pinned memory so that the H2D copy is one DMA;
	// inputs already in page-locked memory (mm_host_alloc) go out directly
	const uint8_t *hc = chunks;
	const int64_t *ho = offs;
	if (n && !(is_pinned_ptr(chunks) && is_pinned_ptr(offs))) {
		const size_t in_bytes = (size_t)n * (4096 + 24);
		if ((rc = ensure_pinned(c, &c->h_in, &c->h_in_bytes, in_bytes))) return rc;
		memcpy(c->h_in, chunks, (size_t)n * 4096);
		memcpy(c->h_in + (size_t)n * 4096, offs, (size_t)n * 24);
		hc = c->h_in;
		ho = (const int64_t *)(c->h_in + (size_t)n * 4096);
	}
	if (c->pipe_min_chunks && n >= c->pipe_min_chunks) {
		bool done = false;
		if ((rc = pipelined_batch_host(c, hc, ho, n, out, &done))) return rc;
		if (done) return MM_OK;
	}
	mm_batch_out bo;
	if ((rc = mm_mesh_batch(c, hc, ho, nullptr, n, &bo))) return rc;
	const size_t bi = (size_t)bo.n_intransparent * 36, bt = (size_t)bo.n_transparent * 36, br = (size_t)n * 16;
	const size_t oi = 0, ot = (bi + 255) & ~(size_t)255, ori = (ot + bt + 255) & ~(size_t)255, ort = ori + ((br + 255) & ~(size_t)255);
	if ((rc = ensure_pinned(c, &c->h_out, &c->h_out_bytes, ort + br + 256))) return rc;
	if ((rc = mm_copy_out(c, (mm_vertex *)(c->h_out + oi), (mm_vertex *)(c->h_out + ot), (mm_range *)(c->h_out + ori),
	                      (mm_range *)(c->h_out + ort))))
		return rc;
	out->intransparent = (const mm_vertex *)(c->h_out + oi);
	out->transparent = (const mm_vertex *)(c->h_out + ot);
	out->ranges_i = (const mm_range *)(c->h_out + ori);
	out->ranges_t = (const mm_range *)(c->h_out + ort);
	out->n_intransparent = bo.n_intransparent;
	out->n_transparent = bo.n_transparent;
	out->n_chunks = n;
	out->path = MM_PATH_PLAIN;
	if (n) {
		c->host_qi_per_chunk = (double)(bo.n_intransparent / 6) / n;
		c->host_qt_per_chunk = (double)(bo.n_transparent / 6) / n;
	}
	return MM_OK;
}

int mm_mesh_for_chunk(mm_ctx *c, const int64_t offs[3], const uint8_t chunk[MM_CHUNK_BYTES], const mm_vertex **vi,
                      uint64_t *ni, const mm_vertex **vt, uint64_t *nt) {
	if (!c || !offs || !chunk || !vi || !ni || !vt || !nt) return MM_ERR_INVALID;
	mm_host_mesh hm;
	const int rc = mm_mesh_batch_host(c, chunk, offs, 1, &hm);
	if (rc) return rc;
	*vi = hm.intransparent; *ni = hm.n_intransparent;
	*vt = hm.transparent; *nt = hm.n_transparent;
	return MM_OK;
}

int mm_debug_trace(mm_ctx *c, uint32_t pairs, uint64_t *out, uint64_t *n_launched) {
	if (!c) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	if (out && c->d_trace) MM_CUDA(c, cudaMemcpy(out, c->d_trace, sizeof(mm::u64) * 8 * c->trace_pairs, cudaMemcpyDeviceToHost));
	if (n_launched) *n_launched = c->trace_next;
	if (pairs != c->trace_pairs) {
		cudaFree(c->d_trace);
		c->d_trace = nullptr;
		c->trace_pairs = 0;
		if (pairs) MM_CUDA(c, cudaMalloc(&c->d_trace, sizeof(mm::u64) * 8 * pairs));
		c->trace_pairs = pairs;
	}
	if (c->d_trace) {  // minima start at all-ones, maxima at zero
		std::vector<mm::u64> init((size_t)8 * pairs);
		for (size_t i = 0; i < init.size(); i++) init[i] = (i & 2) ? 0ull : ~0ull;
		MM_CUDA(c, cudaMemcpy(c->d_trace, init.data(), sizeof(mm::u64) * init.size(), cudaMemcpyHostToDevice));
	}
	c->trace_next = 0;
	return MM_OK;
}

int mm_dev_alloc(mm_ctx *c, size_t bytes, void **d_ptr) {
	if (!c || !d_ptr) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	*d_ptr = nullptr;
	MM_CUDA(c, cudaMalloc(d_ptr, bytes ? bytes : 1));
	return MM_OK;
}
int mm_dev_free(mm_ctx *c, void *d_ptr) {
	if (!c) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	MM_CUDA(c, cudaFree(d_ptr));
	return MM_OK;
}
int mm_dev_upload(mm_ctx *c, void *d_dst, const void *h_src, size_t bytes) {
	if (!c || (bytes && (!d_dst || !h_src))) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, c->stream));
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	return MM_OK;
}
int mm_dev_download(mm_ctx *c, void *h_dst, const void *d_src, size_t bytes) {
	if (!c || (bytes && (!h_dst || !d_src))) return MM_ERR_INVALID;
	if (use_device(c)) return MM_ERR_CUDA;
	MM_CUDA(c, cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, c->stream));
	MM_CUDA(c, cudaStreamSynchronize(c->stream));
	return MM_OK;
}

}  // extern "C"

// ---- mm_mesh_batch_noalloc: the mesher for callers that must not allocate ------------------------------------
// No counterpart in the reference.  The analysis -> expansion
// hand-over (descriptor slots, counts, ticket word) of up to MM_NOALLOC_MAX_CHUNKS chunks lives in a file-scope
// device array instead of cudaMalloc'ed buffers: every GPU that loads the library holds it until exit.  All calls
// share it (contents of the last call stay; the ticket word is cleared on the stream before every pair).
namespace {
constexpr size_t kNaHead = 256;                                                        // ticket word
constexpr size_t kNaCnt = sizeof(uint32_t) * MM_NOALLOC_MAX_CHUNKS;                    // cnt_i, then cnt_t
constexpr size_t kNaDescI = sizeof(uint32_t) * (size_t)MM_NOALLOC_MAX_CHUNKS * mm::kDescStrideI;
constexpr size_t kNaDescT = sizeof(uint32_t) * (size_t)MM_NOALLOC_MAX_CHUNKS * mm::kDescStrideT;
constexpr size_t kNaBytes = kNaHead + 2 * kNaCnt + kNaDescI + kNaDescT;
static_assert(MM_NOALLOC_MAX_CHUNKS % 4 == 0 && MM_NOALLOC_MAX_CHUNKS <= mm::kMaxLaunchChunks, "counts are read four at a time; one launch pair");
static_assert(kNaCnt % 256 == 0 && kNaDescI % 256 == 0, "every part stays 256-byte aligned");
}  // namespace
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];
namespace {
// In host code g_noalloc_scratch is a stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
std::mutex g_noalloc_mutex;
unsigned char *g_noalloc_ptr[64] = {nullptr};
int noalloc_scratch(mm_ctx *c, unsigned char **out) {
	if (c->device < 0 || c->device >= 64) return fail(c, MM_ERR_UNSUPPORTED, "device index %d beyond the table of scratch addresses", c->device);
	std::lock_guard<std::mutex> lock(g_noalloc_mutex);
	if (!g_noalloc_ptr[c->device]) {
		void *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // use_device(c) made c->device current
		if (e != cudaSuccess || !p) return fail(c, MM_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch): %s", cudaGetErrorString(e));
		g_noalloc_ptr[c->device] = (unsigned char *)p;
	}
	*out = g_noalloc_ptr[c->device];
	return MM_OK;
}
}  // namespace

extern "C" {

int mm_mesh_batch_noalloc(mm_ctx *c, const uint8_t *d_chunks, const int64_t *d_offs, uint32_t n, mm_vertex *d_out_i,
                          uint64_t cap_i_vertices, mm_vertex *d_out_t, uint64_t cap_t_vertices, mm_range *d_ranges_i,
                          mm_range *d_ranges_t, uint64_t *d_totals) {
	if (!c || !d_totals || (n && (!d_chunks || !d_offs || !d_ranges_i || !d_ranges_t))) return MM_ERR_INVALID;
	if (n > MM_NOALLOC_MAX_CHUNKS)
		return fail(c, MM_ERR_INVALID, "mm_mesh_batch_noalloc: %u chunks exceed the static scratch of %u chunks (MM_NOALLOC_MAX_CHUNKS); use mm_mesh_batch",
		            n, (unsigned)MM_NOALLOC_MAX_CHUNKS);
	if ((cap_i_vertices && !d_out_i) || (cap_t_vertices && !d_out_t)) return fail(c, MM_ERR_INVALID, "capacity without an output buffer");
	if (n && ((uintptr_t)d_chunks & 15u)) return fail(c, MM_ERR_INVALID, "chunks must be 16-byte aligned");
	if (((uintptr_t)d_offs | (uintptr_t)d_ranges_i | (uintptr_t)d_ranges_t | (uintptr_t)d_totals) & 7u)
		return fail(c, MM_ERR_INVALID, "offs, range records and totals must be 8-byte aligned");
	if (((uintptr_t)d_out_i | (uintptr_t)d_out_t) & 15u) return fail(c, MM_ERR_INVALID, "vertex buffers must be 16-byte aligned");
	if (!c->have_table) return fail(c, MM_ERR_NO_TABLE, "mm_set_block_table has not been called");
	if (use_device(c)) return MM_ERR_CUDA;
	if (n == 0) {
		MM_CUDA(c, cudaMemsetAsync(d_totals, 0, 2 * sizeof(uint64_t), c->stream));
		return MM_OK;
	}
	unsigned char *scr = nullptr;
	int rc;
	if ((rc = noalloc_scratch(c, &scr))) return rc;
	MM_CUDA(c, cudaMemsetAsync(scr, 0, kNaHead, c->stream));  // the ticket word starts every pair at zero

	mm::AnalyseParams a;
	memset(&a, 0, sizeof(a));
	a.chunks = d_chunks;
	a.halo = nullptr;
	a.n = n;
	a.table = c->d_table;
	a.flags = c->d_flags;
	a.cnt_i = reinterpret_cast<uint32_t *>(scr + kNaHead);
	a.cnt_t = reinterpret_cast<uint32_t *>(scr + kNaHead + kNaCnt);
	a.desc_i = reinterpret_cast<uint32_t *>(scr + kNaHead + 2 * kNaCnt);
	a.desc_t = reinterpret_cast<uint32_t *>(scr + kNaHead + 2 * kNaCnt + kNaDescI);
	a.ticket = reinterpret_cast<mm::u64 *>(scr);
	a.ticket_base = 0;
	a.trace = nullptr;
	a.h_done = nullptr;
	a.prev_done = 0;
	// no programmatic dependent launch: the pair runs strictly after whatever is on the stream, and in order
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 0;
	{
		cudaLaunchConfig_t cfg = {};
		cfg.gridDim = dim3(std::min<uint32_t>(n, (uint32_t)(c->sm_count * c->analyse_ctas_per_sm)));
		cfg.blockDim = dim3(mm::kThreads);
		cfg.stream = c->stream;
		cfg.attrs = attr;
		cfg.numAttrs = 1;
		MM_CUDA(c, cudaLaunchKernelEx(&cfg, mm::analyse_kernel<false>, a));
	}
	mm::ExpandParams e;
	memset(&e, 0, sizeof(e));
	e.offs = (const long long *)d_offs;
	e.n = n;
	e.table = c->d_table;
	e.desc_i = a.desc_i; e.desc_t = a.desc_t; e.cnt_i = a.cnt_i; e.cnt_t = a.cnt_t;
	e.out_i = (unsigned char *)d_out_i; e.out_t = (unsigned char *)d_out_t;
	e.cap_i = cap_i_vertices / 6; e.cap_t = cap_t_vertices / 6;
	e.ranges_i = (mm::Range *)d_ranges_i; e.ranges_t = (mm::Range *)d_ranges_t;
	e.carry_in = nullptr;
	e.totals = (mm::u64 *)d_totals;
	e.host_totals = (mm::u64 *)d_totals;  // the kernel writes the two words twice; the caller's device words take both
	e.trace = nullptr;
	e.h_flag = nullptr;
	e.flag_value = 0;
	{
		cudaLaunchConfig_t cfg = {};
		cfg.gridDim = dim3((unsigned)c->expand_grid);
		cfg.blockDim = dim3(mm::kExpandThreads);
		cfg.dynamicSmemBytes = mm::kExpandSmemBytes;
		cfg.stream = c->stream;
		cfg.attrs = attr;
		cfg.numAttrs = 1;
		MM_CUDA(c, cudaLaunchKernelEx(&cfg, mm::expand_kernel, e));
	}
	c->total_launches += 2;
	return MM_OK;
}

}  // extern "C"
