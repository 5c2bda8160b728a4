// mesh_kernel.cuh -- the sm_100a batch mesher, v3: two kernels per batch (device only).
//
// analyse_kernel: persistent CTAs of 128 threads, 8 per SM; every CTA has one chunk in flight, draws
// chunk tickets and runs phases A-D of mesh_core.cuh:
//   A. one bulk async copy (TMA) brings the 4 KiB chunk into shared memory; thread <-> two columns:
//      blocky masks and id-equality masks
//   B. visibility / run starts of the six passes as bit operations on 16-bit z-masks
//   C. 32 + 96 jobs, one per thread: 16x16 bit transposes in registers -> rows in emission order;
//      popc + one warp scan -> quads and non-empty words per (pass, c1) group
//   D. the non-empty words (two rows each) are listed; thread <-> word writes one 32-bit descriptor
//      per quad into the chunk's descriptor slot; plants likewise; the two counts into the count arrays
// The analysis never waits for another chunk: no look-back, no spinning.
//
// expand_kernel: one persistent CTA of 12 warps per SM.  Every CTA scans the per-chunk counts
// into shared memory (exclusive prefix = where each chunk starts in the packed streams, chunk
// order, tightly packed); then each warp takes tiles of 32 consecutive quads of the packed
// stream: quad -> chunk by binary search in shared memory, descriptor -> 216 B per lane into the
// warp's staging tile, one cp.async.bulk (TMA) shared -> global store per tile.  Tiles are
// uniform, so the write-bound part of the job is balanced to one tile per warp.
//
// Consecutive batches chain by programmatic dependent launch: the analysis of batch k+1 runs beside
// the expansion of batch k (see the griddepcontrol calls and DESIGN.md section 3).
// No tensor cores: the path is a scan/compaction bound by the vertex-stream writes.
#pragma once
#include <cuda_runtime.h>
#include "mesh_core.cuh"

namespace mm {

constexpr int kWarpsPerCta = 4;                  // analyse_kernel: one CTA per chunk in flight
constexpr int kThreads = kWarpsPerCta * 32;
#ifndef MM_ANALYSE_CTAS
#define MM_ANALYSE_CTAS 10
#endif
#ifndef MM_EXPAND_WARPS
#define MM_EXPAND_WARPS 12
#endif
#ifndef MM_EXPAND_MINBLOCKS
#define MM_EXPAND_MINBLOCKS 3
#endif
#ifndef MM_ANALYSE_UNIFORM_EXIT
#define MM_ANALYSE_UNIFORM_EXIT 1
#endif
#ifndef MM_ANALYSE_PREFETCH
#define MM_ANALYSE_PREFETCH 1
#endif
#ifndef MM_EXPAND_PREFETCH
#define MM_EXPAND_PREFETCH 0
#endif
#ifndef MM_EXPAND_WARP_SEARCH
#define MM_EXPAND_WARP_SEARCH 1
#endif
constexpr int kAnalyseCtasPerSm = MM_ANALYSE_CTAS;   // register budget of the analysis: 65536 / (128 * this)
static_assert(kThreads == 128, "phase C hands one of 32 + 96 jobs to every thread");
constexpr int kExpandWarps = MM_EXPAND_WARPS;    // expand_kernel
constexpr int kExpandThreads = kExpandWarps * 32;
constexpr uint32_t kMaxLaunchChunks = 4096;      // the prefix of one launch lives in shared memory
// expand: table | prefix_i[4096+4] | prefix_t[4096+4] | warp totals | stage tiles
constexpr size_t kExpandPrefixWords = kMaxLaunchChunks + 4;
constexpr size_t kExpandSmemBytes = 2048 + 2 * 4 * kExpandPrefixWords + 512 + kExpandWarps * (size_t)(kStageBytes + 16);

struct ExpandParams {
	const long long *offs;        // n * 3
	uint32_t n;
	const TabEntry *table;
	const uint32_t *desc_i, *desc_t, *cnt_i, *cnt_t;
	unsigned char *out_i;         // intransparent vertex stream
	unsigned char *out_t;         // transparent vertex stream
	u64 cap_i, cap_t;             // arena capacities in quads (216 B units)
	Range *ranges_i, *ranges_t;   // n entries each
	const u64 *carry_in;          // quads already in the streams (earlier launches of the batch) or NULL
	u64 *totals;                  // [0] quads in out_i, [1] quad slots in out_t after this launch
	u64 *host_totals;             // the same two words in mapped pinned host memory
	u64 *trace;                   // NULL, or 4 words of this launch: first CTA start, first CTA past the wait, -, last CTA end (ns)
	// FUSED single-chunk launch over mapped host memory only: a mapped host word that receives flag_value once the
	// vertices have landed in host memory, so that the caller polls a cache line instead of synchronising the stream
	u64 *h_flag;
	u64 flag_value;
};

struct AnalyseParams {
	const unsigned char *chunks;  // n * 4096
	const unsigned char *halo;    // halo mode: n * 1536 (see halo_row in mesh_core.cuh), 16-byte aligned; else unused
	uint32_t n;                   // <= kMaxLaunchChunks
	const TabEntry *table;        // 256 entries
	const unsigned char *flags;   // 256 flag bytes (kFlagBlocky / kFlagCrossed)
	uint32_t *desc_i;             // n * kDescStrideI descriptor words
	uint32_t *desc_t;             // n * kDescStrideT
	uint32_t *cnt_i;              // quads per chunk
	uint32_t *cnt_t;              // transparent quad slots per chunk (4 per plant)
	u64 *ticket;                  // monotonically increasing across launches
	u64 ticket_base;              // its value at launch
	u64 *trace;                   // NULL, or 4 words of this launch: first CTA start, -, last CTA work done, last CTA end (ns)
	// Completion signal of the batch BEFORE this launch: once this kernel is past its dependency wait, everything
	// launched earlier on the stream has completed, and one thread says so to the host (mm_batch_wait polls the
	// word) -- a batch can be handed to a consumer while later ones still run, at no cost to the batch itself.
	u64 *h_done;                  // NULL, or mapped pinned host word
	u64 prev_done;                // what to write there
	ExpandParams x;               // FUSED instantiation only: where the one chunk's vertices go
};


__device__ __forceinline__ Col lds_col(const unsigned char *p) {
	const uint4 v = *reinterpret_cast<const uint4 *>(p);
	Col c;
	c.w[0] = v.x; c.w[1] = v.y; c.w[2] = v.z; c.w[3] = v.w;
	return c;
}

// ---- TMA bulk load of one chunk (4 KiB, global -> shared) ------------------------------
__device__ __forceinline__ void mbar_init(unsigned long long *mbar) {
	const uint32_t a = (uint32_t)__cvta_generic_to_shared(mbar);
	asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(a) : "memory");
}
__device__ __forceinline__ void bulk_load_g2s(void *sdst, const void *gsrc, uint32_t bytes, unsigned long long *mbar) {
	const uint32_t d = (uint32_t)__cvta_generic_to_shared(sdst), b = (uint32_t)__cvta_generic_to_shared(mbar);
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d),
	             "l"(gsrc), "r"(bytes), "r"(b)
	             : "memory");
}
// pulls a chunk into L2 ahead of its bulk load (no destination, no completion to wait for)
__device__ __forceinline__ void bulk_prefetch_l2(const void *gsrc, uint32_t bytes) {
	asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *mbar, uint32_t parity) {
	const uint32_t b = (uint32_t)__cvta_generic_to_shared(mbar);
	uint32_t done;
	do {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
		             : "=r"(done)
		             : "r"(b), "r"(parity)
		             : "memory");
	} while (!done);
}
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
		if (lane >= d) v += o;
	}
	return v;
}

// ---- TMA bulk store of a staging tile ------------------------------------------------
__device__ __forceinline__ void fence_proxy_async_smem() {
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_s2g(void *gdst, const void *ssrc, uint32_t bytes) {
	const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(ssrc);
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(saddr),
	             "r"(bytes)
	             : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() {
	asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// programmatic dependent launch (sm_90+): let the next kernel of the stream start its
// prologue early / wait until everything the previous kernel wrote is visible
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// timeline of a launch for mm_debug_trace (one thread per CTA; words 0,1 keep the minimum, 2,3 the maximum)
__device__ __forceinline__ void trace_mark(u64 *trace, int word) {
	if (trace == nullptr) return;
	u64 t;
	asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
	if (word < 2) atomicMin(trace + word, t);
	else atomicMax(trace + word, t);
}

// Called by ONE lane after the tile is complete and fenced.  `g` is the byte address
// of the tile in the stream (8-byte aligned); `sp` is the tile in shared memory with
// sp == g (mod 16), so the 16-byte aligned middle goes out as one bulk copy and at
// most 8 bytes at either end as ordinary stores.
__device__ __forceinline__ void store_tile(unsigned char *g, const unsigned char *sp, uint32_t bytes) {
	uint32_t head = (uint32_t)((16u - (uint32_t)((uintptr_t)g & 15u)) & 15u);
	if (head > bytes) head = bytes;
	if (head) *reinterpret_cast<uint2 *>(g) = *reinterpret_cast<const uint2 *>(sp);
	const uint32_t body = (bytes - head) & ~15u;
	if (body) {
		bulk_store_s2g(g + head, sp + head, body);
		bulk_commit();
	}
	if (bytes - head - body)
		*reinterpret_cast<uint2 *>(g + head + body) = *reinterpret_cast<const uint2 *>(sp + head + body);
}

// ---- kernel 1: analysis ------------------------------------------------------------------
// One CTA of kThreads threads per chunk in flight; CTAs are persistent and draw chunk tickets.
// HALO = the optional mode in which the touching layers of the six neighbour chunks occlude faces (an
// extension; false = the reference's semantics, lib.rs:277-280, and the instantiation every number in
// DESIGN.md is measured on).
// FUSED = the single-chunk form (the reference's own call pattern, one chunk per call, client.rs:259-267): a grid
// of one CTA analyses the chunk and expands it right away -- one launch instead of two and no hand-over.
template <bool HALO, bool FUSED = false>
__global__ void __launch_bounds__(kThreads, kAnalyseCtasPerSm) analyse_kernel(const AnalyseParams p) {
	__shared__ ChunkSmem s;
	__shared__ __align__(16) unsigned char flags[256];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const unsigned full = 0xffffffffu;

	// the expand kernel may set itself up while this one works.  Not so in the FUSED form: it owns no expand
	// kernel whose wait would hold the next launch back until the hand-over buffers are free again, so it lets
	// its dependents go only once everything before it has completed (below, after its own wait).
	if (!FUSED) pdl_launch_dependents();
	if (tid == 0) trace_mark(p.trace, 0);
	if (tid < 16) reinterpret_cast<uint4 *>(flags)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.flags) + tid);
	if (warp == 1) init_warp_smem(s, lane);
	if (tid == 0) {
		mbar_init(&s.mbar);
		fence_proxy_async_smem();  // the barrier is visible to the bulk-copy engine
		// FUSED: one CTA, one chunk -- no ticket is drawn, so a following launch may draw its own at once
		s.ticket[0] = FUSED ? 0u : (uint32_t)min(atomicAdd(p.ticket, 1ull) - p.ticket_base, (u64)0xffffffffu);
		s.amb[0] = 0;
	}
	unsigned char *raw = smem_raw(s);
	const TabEntry *tab = p.table;  // only read for chunks with ambiguous ids (fix_equal)
	uint32_t par = 0;
	for (;; par ^= 1u) {
		__syncthreads();  // the previous chunk is done with shared memory; this round's ticket is visible
		const uint32_t c = s.ticket[par];
		if (c >= p.n) break;
		if (tid == 0) {
			bulk_load_g2s(raw, p.chunks + (size_t)c * 4096u, 4096u, &s.mbar);  // TMA: 4 KiB chunk -> shared memory
			// the next ticket travels while this chunk is analysed
			const uint32_t nxt = FUSED ? 0xffffffffu : (uint32_t)min(atomicAdd(p.ticket, 1ull) - p.ticket_base, (u64)0xffffffffu);
			s.ticket[par ^ 1u] = nxt;
			s.amb[par ^ 1u] = 0;
#if MM_ANALYSE_PREFETCH
			if (nxt < p.n) bulk_prefetch_l2(p.chunks + (size_t)nxt * 4096u, 4096u);  // and its chunk comes as far as L2
#endif
		}
		Col hrow;
		if (HALO && tid < 96) {  // one 16-byte row of the halo slabs per thread, in flight beside the chunk
			const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p.halo + (size_t)c * 1536u) + tid);
			hrow.w[0] = v.x; hrow.w[1] = v.y; hrow.w[2] = v.z; hrow.w[3] = v.w;
		}
		mbar_wait(&s.mbar, par);

		// A. thread <-> columns col = tid, tid + 128 (col = x*16 + y): blocky mask over z per column
		uint32_t crossed_cols = 0, amb = 0;
		Col own[256 / kThreads];
		bool uniform = true;
#pragma unroll
		for (int j = 0; j < 256 / kThreads; j++) {
			own[j] = load_col(raw + (kThreads * j + tid) * 16);
			const uint32_t w0 = own[j].w[0];
			uniform = uniform && w0 == own[j].w[1] && w0 == own[j].w[2] && w0 == own[j].w[3] &&
			          w0 == __byte_perm(w0, 0, 0) && w0 == *reinterpret_cast<const uint32_t *>(raw);
		}
		// a chunk of one undrawn id (all air: a quarter of a terrain region) is done here
		if (MM_ANALYSE_UNIFORM_EXIT && __syncthreads_and(uniform) && (flags[raw[0]] & (kFlagBlocky | kFlagCrossed)) == 0) {
			if (tid == 0) {
				p.cnt_i[c] = 0;
				p.cnt_t[c] = 0;
			}
			continue;
		}
		if (HALO && tid < 96) halo_row(s, flags, hrow, tid);
#pragma unroll
		for (int j = 0; j < 256 / kThreads; j++) {
			const int col = kThreads * j + tid;
			uint32_t orf;
			const uint32_t bz = col_flags(flags, own[j], orf);
			smem_Bzp(s)[((col >> 4) + 1) * kBzStride + (col & 15) + 1] = (uint16_t)bz;
			crossed_cols |= ((orf >> 1) & 1u) << j;
			amb |= orf;
		}
		amb = __reduce_or_sync(full, amb);
		if (lane == 0 && amb) atomicOr(&s.amb[par], amb);
		__syncthreads();
		amb = s.amb[par];
		if (amb & kFlagBlocky) {  // else: no cube in the chunk (e.g. all air)
			uint32_t *di = p.desc_i + (size_t)c * kDescStrideI;
			// B. run starts over z for the six passes
#pragma unroll 1
			for (int j = 0; j < 256 / kThreads; j++) {
				const int col = kThreads * j + tid;
				col_starts<HALO>(s, tab, raw, col >> 4, col & 15, amb);
			}
			__syncthreads();
			// C. masks over the merge axes: 32 blocky jobs + 96 groups = one job per thread
			if (tid < 32) transpose_blocky(s, tid);
			else s.gcnt[tid - 32] = transpose_group(s, tid - 32);
			if (HALO && tid >= 124) transpose_halo(s, tid - 124);  // Bzp's border is read-only here
			__syncthreads();
			// exclusive scan of quads (low half) and non-empty words (high half) over the 96 groups, by every warp
			// for itself (no hand-over through shared memory, one barrier less): the thread of group
			// m = tid - 32 = (warp - 1) * 32 + lane finds its own prefix in round warp - 1
			uint32_t run = 0, mine = 0;
#pragma unroll
			for (int k = 0; k < 3; k++) {
				const uint32_t v = s.gcnt[lane + 32 * k];
				const uint32_t inc = warp_incl_scan(v, lane);
				if (k == warp - 1) mine = run + inc - v;
				run += __shfl_sync(full, inc, 31);
			}
			if (tid >= 32) list_group_words(s, tid - 32, mine);
			__syncthreads();
			// D. descriptors: thread <-> non-empty word (two rows)
			const uint32_t nwords = run >> 16;
#pragma unroll 1
			for (uint32_t i = (uint32_t)tid; i < nwords; i += kThreads) {
				const uint32_t e = smem_wl(s)[i];
				const uint32_t widx = e & 1023u;
				emit_word_descs<HALO>(s, widx, reinterpret_cast<const uint32_t *>(s.S)[widx], di + (e >> 10));
			}
			if (tid == 0) p.cnt_i[c] = run & 0xffffu;
		} else if (tid == 0) {
			p.cnt_i[c] = 0;
		}
		if (amb & kFlagCrossed) {
			__syncthreads();  // wl is dead: Cz takes its place
			uint16_t *Cz = smem_Cz(s);
#pragma unroll
			for (int j = 0; j < 256 / kThreads; j++) {
				const int col = kThreads * j + tid;
				uint32_t cz = 0;
				if ((crossed_cols >> j) & 1u) cz = col_crossed(flags, lds_col(raw + col * 16));
				Cz[col] = (uint16_t)cz;
			}
			__syncthreads();
			if (warp == 0) {
				const uint32_t mine = count_plants8(s, lane);
				const uint32_t inc = warp_incl_scan(mine, lane);
				if (mine) emit_plant_descs(s, lane, p.desc_t + (size_t)c * kDescStrideT + (inc - mine));
				if (lane == 31) p.cnt_t[c] = inc * (uint32_t)kPlantQuads;
			}
		} else if (tid == 0) {
			p.cnt_t[c] = 0;
		}
	}
	// Programmatic dependent launch: this grid runs beside the expansion of the previous batch; it must
	// not complete (and so release the expansion of this batch onto the arenas) before that one has.
	if (tid == 0) trace_mark(p.trace, 2);
	pdl_wait();
	if (tid == 0) trace_mark(p.trace, 3);
	if (p.h_done != nullptr && blockIdx.x == 0 && tid == 0) *reinterpret_cast<volatile u64 *>(p.h_done) = p.prev_done;
	if (FUSED) pdl_launch_dependents();
	if (FUSED) {  // p.n == 1, one CTA: the chunk's descriptors and counts are this CTA's own writes
		__shared__ __align__(16) unsigned char stage_all[kWarpsPerCta * (kStageBytes + 16)];
		__syncthreads();
		const ExpandParams &x = p.x;
		const uint32_t nq = p.cnt_i[0], nt = p.cnt_t[0];
		if (tid == 0) {
			Range ri, rt;
			ri.first_vertex = 0; ri.vertex_count = nq * 6u; ri.reserved = 0;
			rt.first_vertex = 0; rt.vertex_count = nt * 6u; rt.reserved = 0;
			x.ranges_i[0] = ri;
			x.ranges_t[0] = rt;
			x.totals[0] = nq; x.totals[1] = nt;
			x.host_totals[0] = nq; x.host_totals[1] = nt;
			__threadfence_system();
		}
		const uint32_t tiles_i = nq <= x.cap_i ? (nq + kTileQuads - 1) / kTileQuads : 0u;
		const uint32_t tiles_t = nt <= x.cap_t ? (nt + kTileQuads - 1) / kTileQuads : 0u;
		const long long ox = __ldg(x.offs), oy = __ldg(x.offs + 1), oz = __ldg(x.offs + 2);
		unsigned char *stage = stage_all + (size_t)warp * (kStageBytes + 16);
#pragma unroll 1
		for (uint32_t t = (uint32_t)warp; t < tiles_i + tiles_t; t += kWarpsPerCta) {
			const bool face = t < tiles_i;
			const uint32_t q0 = (face ? t : t - tiles_i) * kTileQuads;
			const uint32_t tn = min((uint32_t)kTileQuads, (face ? nq : nt) - q0);
			unsigned char *g = face ? x.out_i + (u64)q0 * kQuadBytes : x.out_t + (u64)q0 * kQuadBytes;
			unsigned char *sp = stage + ((uintptr_t)g & 8u);
			if (lane == 0) bulk_wait_read_all();
			__syncwarp();
			if ((uint32_t)lane < tn) {
				const uint32_t q = q0 + (uint32_t)lane;
				if (face) expand_face_desc(tab, ox, oy, oz, p.desc_i[q], sp + lane * kQuadBytes);
				else expand_plant_desc(tab, ox, oy, oz, p.desc_t[q >> 2], lane & 3, sp + lane * kQuadBytes);
			}
			fence_proxy_async_smem();
			__syncwarp();
			if (lane == 0) store_tile(g, sp, tn * kQuadBytes);
		}
		if (lane == 0) bulk_wait_all();
		if (x.h_flag != nullptr) {
			__syncthreads();
			if (tid == 0) {
				__threadfence_system();
				*reinterpret_cast<volatile u64 *>(x.h_flag) = x.flag_value;
			}
		}
	}
}

// ---- kernel 2: expansion -------------------------------------------------------------------
// Exclusive prefixes of cnt_i / cnt_t [0..n) into pre_i / pre_t [0..lim), lim = the span the binary
// search can touch; entries at and beyond n hold 0xffffffff so that a fixed-depth search never
// steps past n.  Warp w scans the contiguous span [w * rounds * 128, (w+1) * rounds * 128), four
// entries per lane and round (16-byte loads and stores, no bank conflicts); the warp totals are
// combined through shared memory.  Returns the two totals.  All kExpandThreads threads take part.
constexpr int kMaxScanRounds = (kMaxLaunchChunks + kExpandWarps * 128 - 1) / (kExpandWarps * 128);
__device__ __forceinline__ void block_scan_counts(const uint32_t *cnt_i, const uint32_t *cnt_t, uint32_t n, uint32_t lim,
                                                  uint32_t *pre_i, uint32_t *pre_t, uint32_t *warp_tot, int tid,
                                                  uint32_t &tot_i, uint32_t &tot_t) {
	const int lane = tid & 31, warp = tid >> 5;
	const uint32_t rounds = (lim + kExpandWarps * 128 - 1) / (kExpandWarps * 128);
	const uint32_t first = (uint32_t)warp * rounds * 128u + (uint32_t)lane * 4u;
	uint4 vi[kMaxScanRounds], vt[kMaxScanRounds];
	uint32_t ei[kMaxScanRounds], et[kMaxScanRounds];  // exclusive prefix of this lane's four, relative to the warp's span
	uint32_t run_i = 0, run_t = 0;
#pragma unroll
	for (int r = 0; r < kMaxScanRounds; r++) {
		vi[r] = make_uint4(0, 0, 0, 0);
		vt[r] = make_uint4(0, 0, 0, 0);
		const uint32_t i = first + (uint32_t)r * 128u;
		if ((uint32_t)r < rounds && i < n) {  // the count arrays are padded to kMaxLaunchChunks
			vi[r] = *reinterpret_cast<const uint4 *>(cnt_i + i);
			vt[r] = *reinterpret_cast<const uint4 *>(cnt_t + i);
			if (i + 1 >= n) vi[r].y = vt[r].y = 0;
			if (i + 2 >= n) vi[r].z = vt[r].z = 0;
			if (i + 3 >= n) vi[r].w = vt[r].w = 0;
		}
	}
#pragma unroll
	for (int r = 0; r < kMaxScanRounds; r++) {
		if ((uint32_t)r < rounds) {
			const uint32_t si = vi[r].x + vi[r].y + vi[r].z + vi[r].w, st = vt[r].x + vt[r].y + vt[r].z + vt[r].w;
			const uint32_t ii = warp_incl_scan(si, lane), it = warp_incl_scan(st, lane);
			ei[r] = run_i + ii - si;
			et[r] = run_t + it - st;
			run_i += __shfl_sync(0xffffffffu, ii, 31);
			run_t += __shfl_sync(0xffffffffu, it, 31);
		}
	}
	if (lane == 0) {
		warp_tot[warp] = run_i;
		warp_tot[16 + warp] = run_t;
	}
	__syncthreads();
	static_assert(kExpandWarps <= 16, "one lane per warp total, two streams in one warp scan");
	{
		// lanes 0..15: stream i, lanes 16..31: stream t; warps that do not exist count zero
		const uint32_t wt = (lane & 15) < kExpandWarps ? warp_tot[lane] : 0u;
		uint32_t inc = wt;
#pragma unroll
		for (int d = 1; d < 16; d <<= 1) {
			const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
			if ((lane & 15) >= d) inc += o;
		}
		const uint32_t ex = inc - wt;
		run_i = __shfl_sync(0xffffffffu, ex, warp);
		run_t = __shfl_sync(0xffffffffu, ex, 16 + warp);
		tot_i = __shfl_sync(0xffffffffu, inc, 15);
		tot_t = __shfl_sync(0xffffffffu, inc, 31);
	}
#pragma unroll
	for (int r = 0; r < kMaxScanRounds; r++) {
		const uint32_t i = first + (uint32_t)r * 128u;
		if ((uint32_t)r < rounds && i < lim) {
			uint4 a, b;
			a.x = run_i + ei[r]; a.y = a.x + vi[r].x; a.z = a.y + vi[r].y; a.w = a.z + vi[r].z;
			b.x = run_t + et[r]; b.y = b.x + vt[r].x; b.z = b.y + vt[r].y; b.w = b.z + vt[r].z;
			if (i >= n) a.x = b.x = 0xffffffffu;
			if (i + 1 >= n) a.y = b.y = 0xffffffffu;
			if (i + 2 >= n) a.z = b.z = 0xffffffffu;
			if (i + 3 >= n) a.w = b.w = 0xffffffffu;
			*reinterpret_cast<uint4 *>(pre_i + i) = a;
			*reinterpret_cast<uint4 *>(pre_t + i) = b;
		}
	}
	__syncthreads();
}

// largest c in [0, n) with pre[c] <= q  (q < pre[n]).  pre[] is filled up to kMaxLaunchChunks with
// 0xffffffff beyond n, so a search of fixed depth needs no bounds: 12 steps of load, compare, add.
__device__ __forceinline__ uint32_t find_chunk(const uint32_t *pre, uint32_t q) {
	static_assert(kMaxLaunchChunks == 4096, "12 fixed steps");
	uint32_t lo = 0;
#pragma unroll
	for (uint32_t step = kMaxLaunchChunks / 2; step >= 1; step >>= 1)
		if (pre[lo + step] <= q) lo += step;
	return lo;
}

// The chunk of the first quad of a tile, found by the whole warp: three rounds of 32 / 32 / 4 probes (128-, 4-, 1-entry
// strides over the 4096 prefixes) instead of twelve dependent steps per lane.  Same for all lanes.
// `top` = every 128th prefix, packed (the strided probes themselves would all hit one bank).
__device__ __forceinline__ uint32_t find_chunk_warp(const uint32_t *pre, const uint32_t *top, uint32_t q0, int lane) {
	static_assert(kMaxLaunchChunks == 4096, "32 x 32 x 4 probes");
	const unsigned full = 0xffffffffu;
	const uint32_t b1 = (uint32_t)__popc(__ballot_sync(full, top[lane] <= q0)) - 1u;     // pre[0] = 0 <= q0: at least one
	const uint32_t b2 = (uint32_t)__popc(__ballot_sync(full, pre[b1 * 128u + lane * 4] <= q0)) - 1u;
	const uint32_t base = b1 * 128u + b2 * 4u;
	const uint32_t b3 = (uint32_t)__popc(__ballot_sync(full, lane < 4 && pre[base + (lane & 3)] <= q0)) - 1u;
	return base + b3;
}

__global__ void __launch_bounds__(kExpandThreads, MM_EXPAND_MINBLOCKS) expand_kernel(const ExpandParams p) {
	extern __shared__ __align__(128) unsigned char smem_raw_bytes[];
	TabEntry *tab = reinterpret_cast<TabEntry *>(smem_raw_bytes);
	uint32_t *pre_i = reinterpret_cast<uint32_t *>(smem_raw_bytes + 2048);
	uint32_t *pre_t = pre_i + kExpandPrefixWords;
	uint32_t *warp_tot = pre_t + kExpandPrefixWords;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	uint32_t *top_i = warp_tot + 64, *top_t = top_i + 32;  // every 128th prefix, for the first round of find_chunk_warp
	unsigned char *stage = smem_raw_bytes + 2048 + 2 * 4 * kExpandPrefixWords + 512 + (size_t)warp * (kStageBytes + 16);

	if (tid < 128) reinterpret_cast<uint4 *>(tab)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.table) + tid);
	const uint32_t lim = kMaxLaunchChunks;  // find_chunk() may touch every entry below that
	// Programmatic dependent launch.  The analysis of this batch has completed once the wait returns,
	// and with it (its CTAs end in a wait of their own) the expansion of the previous batch: the
	// arenas and the other half of the hand-over buffers are free.  Only then may the next analysis
	// start: it overlaps the tile loop below.
	if (tid == 0) trace_mark(p.trace, 0);
	pdl_wait();
	pdl_launch_dependents();
	if (tid == 0) trace_mark(p.trace, 1);
	uint32_t tot_i, tot_t;
	block_scan_counts(p.cnt_i, p.cnt_t, p.n, lim, pre_i, pre_t, warp_tot, tid, tot_i, tot_t);
	if (tid < 64) (tid < 32 ? top_i : top_t - 32)[tid] = (tid < 32 ? pre_i : pre_t)[(tid & 31) * 128];
	__syncthreads();
	u64 carry_i = 0, carry_t = 0;
	if (p.carry_in) {
		carry_i = p.carry_in[0];
		carry_t = p.carry_in[1];
	}
	// per-chunk ranges and the totals: spread over the grid
	for (uint32_t c = blockIdx.x * kExpandThreads + tid; c < p.n; c += gridDim.x * kExpandThreads) {
		const uint32_t ni = (c + 1 < p.n ? pre_i[c + 1] : tot_i) - pre_i[c], nt = (c + 1 < p.n ? pre_t[c + 1] : tot_t) - pre_t[c];
		Range ri, rt;
		ri.first_vertex = (carry_i + pre_i[c]) * 6ull; ri.vertex_count = ni * 6u; ri.reserved = 0;
		rt.first_vertex = (carry_t + pre_t[c]) * 6ull; rt.vertex_count = nt * 6u; rt.reserved = 0;
		p.ranges_i[c] = ri;
		p.ranges_t[c] = rt;
	}
	if (blockIdx.x == 0 && tid == 0) {
		p.totals[0] = carry_i + tot_i;
		p.totals[1] = carry_t + tot_t;
		p.host_totals[0] = carry_i + tot_i;
		p.host_totals[1] = carry_t + tot_t;
		__threadfence_system();
	}

	const bool fits_i = carry_i + tot_i <= p.cap_i, fits_t = carry_t + tot_t <= p.cap_t;
	const uint32_t tiles_i = fits_i ? (tot_i + kTileQuads - 1) / kTileQuads : 0u;
	const uint32_t tiles_t = fits_t ? (tot_t + kTileQuads - 1) / kTileQuads : 0u;
	const uint32_t gw = blockIdx.x * kExpandWarps + warp, nw = gridDim.x * kExpandWarps;
	const uint32_t tiles = tiles_i + tiles_t;
	// what a lane needs for its quad of a tile; fetched one tile ahead so that the shared-memory search
	// and the L2 round trip of tile k+1 hide behind the expansion of tile k
	struct Fetched {
		uint32_t d;
		long long ox, oy, oz;
	};
	auto fetch = [&](uint32_t t, Fetched &f) {
		const bool face = t < tiles_i;
		const uint32_t q0 = (face ? t : t - tiles_i) * kTileQuads;
		const uint32_t q = q0 + (uint32_t)lane;
		const uint32_t *pre = face ? pre_i : pre_t;
		// most tiles lie inside one chunk: one search by the warp; only a tile that crosses into the next chunk
		// (pre[c0 + 1] inside it) sends its lanes searching on their own
		const uint32_t c0 = MM_EXPAND_WARP_SEARCH ? find_chunk_warp(pre, face ? top_i : top_t, q0, lane) : 0u;
		const bool one = MM_EXPAND_WARP_SEARCH && pre[c0 + 1] >= q0 + (uint32_t)kTileQuads;
		f.d = 0; f.ox = 0; f.oy = 0; f.oz = 0;
		if (q < (face ? tot_i : tot_t)) {
			const uint32_t c = one ? c0 : find_chunk(pre, q);
			const uint32_t r = q - pre[c];
			f.d = face ? __ldg(p.desc_i + (size_t)c * kDescStrideI + r) : __ldg(p.desc_t + (size_t)c * kDescStrideT + (r >> 2));
			f.ox = __ldg(p.offs + 3 * (size_t)c);
			f.oy = __ldg(p.offs + 3 * (size_t)c + 1);
			f.oz = __ldg(p.offs + 3 * (size_t)c + 2);
		}
	};
	Fetched cur, nxt;
	if (MM_EXPAND_PREFETCH && gw < tiles) fetch(gw, cur);
#pragma unroll 1
	for (uint32_t t = gw; t < tiles; t += nw) {
		if (!MM_EXPAND_PREFETCH) fetch(t, cur);
		else if (t + nw < tiles) fetch(t + nw, nxt);
		const bool face = t < tiles_i;
		const uint32_t q0 = (face ? t : t - tiles_i) * kTileQuads;
		const uint32_t tn = min((uint32_t)kTileQuads, (face ? tot_i : tot_t) - q0);
		const bool act = (uint32_t)lane < tn;
		unsigned char *g = (face ? p.out_i + (carry_i + q0) * (u64)kQuadBytes : p.out_t + (carry_t + q0) * (u64)kQuadBytes);
		unsigned char *sp = stage + ((uintptr_t)g & 8u);
		if (lane == 0) bulk_wait_read_all();  // the staging tile is free once the last bulk store has read it
		__syncwarp();
		if (act) {
			if (face) expand_face_desc(tab, cur.ox, cur.oy, cur.oz, cur.d, sp + lane * kQuadBytes);
			else expand_plant_desc(tab, cur.ox, cur.oy, cur.oz, cur.d, lane & 3, sp + lane * kQuadBytes);
		}
		fence_proxy_async_smem();
		__syncwarp();
		if (lane == 0) store_tile(g, sp, tn * kQuadBytes);
		if (MM_EXPAND_PREFETCH) cur = nxt;
	}
	if (lane == 0) bulk_wait_all();  // all bulk stores of this warp have landed before it exits
	if (tid == 0) trace_mark(p.trace, 3);
}

}  // namespace mm
