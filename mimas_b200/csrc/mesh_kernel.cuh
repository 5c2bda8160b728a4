// mesh_kernel.cuh -- the sm_100a batch mesher, v3: two kernels per batch (device only).
//
// analyse_kernel: persistent CTAs of 8 warps; every WARP is an independent chunk worker that
// draws chunk tickets and runs phases A-D of mesh_core.cuh with its 32 lanes:
//   A. one bulk async copy (TMA) brings the 4 KiB chunk into shared memory; blocky masks and
//      id-equality masks per column
//   B. visibility / run starts of the six passes as bit operations on 16-bit z-masks
//   C. popc -> quads per (pass, c1) group, warp shuffle scan -> offsets inside the chunk
//   D. in-register 16x16 bit transposes -> rows in emission order; one 32-bit descriptor per
//      quad / per plant into the chunk's descriptor slot, the two counts into the count arrays
// The analysis never waits for another chunk: no look-back, no spinning.
//
// expand_kernel: one persistent CTA of 16 warps per SM.  Every CTA scans the per-chunk counts
// into shared memory (exclusive prefix = where each chunk starts in the packed streams, chunk
// order, tightly packed); then each warp takes tiles of 32 consecutive quads of the packed
// stream: quad -> chunk by binary search in shared memory, descriptor -> 216 B per lane into the
// warp's staging tile, one cp.async.bulk (TMA) shared -> global store per tile.  Tiles are
// uniform, so the write-bound part of the job is balanced to one tile per warp.
// No tensor cores: the path is a scan/compaction bound by the vertex-stream writes.
#pragma once
#include <cuda_runtime.h>
#include "mesh_core.cuh"

namespace mm {

constexpr int kWarpsPerCta = 8;                  // analyse_kernel
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kExpandWarps = 16;                 // expand_kernel
constexpr int kExpandThreads = kExpandWarps * 32;
constexpr uint32_t kMaxLaunchChunks = 8192;      // the prefix of one launch lives in shared memory
constexpr int kScanPerThread = kMaxLaunchChunks / kExpandThreads;  // 16
constexpr size_t kAnalyseTableBytes = 2048 + 256;  // block table + flag bytes
constexpr size_t kAnalyseSmemBytes = kAnalyseTableBytes + kWarpsPerCta * sizeof(WarpSmem);
// expand: table | prefix_i[8192+4] | prefix_t[8192+4] | warp totals | stage tiles
constexpr size_t kExpandPrefixWords = kMaxLaunchChunks + 4;
constexpr size_t kExpandSmemBytes = 2048 + 2 * 4 * kExpandPrefixWords + 256 + kExpandWarps * (size_t)(kStageBytes + 16);

struct AnalyseParams {
	const unsigned char *chunks;  // n * 4096
	uint32_t n;                   // <= kMaxLaunchChunks
	const TabEntry *table;        // 256 entries
	const unsigned char *flags;   // 256 flag bytes (kFlagBlocky / kFlagCrossed)
	uint32_t *desc_i;             // n * kDescStrideI descriptor words
	uint32_t *desc_t;             // n * kDescStrideT
	uint32_t *cnt_i;              // quads per chunk
	uint32_t *cnt_t;              // transparent quad slots per chunk (4 per plant)
	u64 *ticket;                  // monotonically increasing across launches
	u64 ticket_base;              // its value at launch
};

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
__device__ __forceinline__ uint32_t warp_sum(uint32_t v) { return __reduce_add_sync(0xffffffffu, v); }
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
__global__ void __launch_bounds__(kThreads, 2) analyse_kernel(const AnalyseParams p) {
	extern __shared__ __align__(128) unsigned char smem_raw_bytes[];
	TabEntry *tab = reinterpret_cast<TabEntry *>(smem_raw_bytes);
	unsigned char *flags = smem_raw_bytes + 2048;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	WarpSmem &s = *reinterpret_cast<WarpSmem *>(smem_raw_bytes + kAnalyseTableBytes + (size_t)warp * sizeof(WarpSmem));
	const unsigned full = 0xffffffffu;

	pdl_launch_dependents();  // the expand kernel may set itself up while this one works
	if (tid < 128) reinterpret_cast<uint4 *>(tab)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.table) + tid);
	if (tid < 16) reinterpret_cast<uint4 *>(flags)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.flags) + tid);
	init_warp_smem(s, lane);
	if (lane == 0) {
		mbar_init(&s.mbar);
		fence_proxy_async_smem();  // the barrier is visible to the bulk-copy engine
	}
	__syncthreads();

	const int h = lane >> 4, y = lane & 15;
	unsigned char *raw = smem_raw(s);
	uint32_t parity = 0;
	u64 next = 0;
	if (lane == 0) next = atomicAdd(p.ticket, 1ull) - p.ticket_base;
	for (;;) {
		const u64 ticket = __shfl_sync(full, next, 0);  // also orders the previous chunk's smem reads before this one's writes
		if (ticket >= (u64)p.n) break;
		const uint32_t c = (uint32_t)ticket;
		const unsigned char *chunk = p.chunks + (size_t)c * 4096u;
		if (lane == 0) {
			bulk_load_g2s(raw, chunk, 4096u, &s.mbar);            // TMA: 4 KiB chunk -> shared memory
			next = atomicAdd(p.ticket, 1ull) - p.ticket_base;     // the next ticket travels while this chunk is analysed
		}
		mbar_wait(&s.mbar, parity);
		parity ^= 1u;

		// A. lane (h, y) owns columns (x = 2j + h, y), j = 0..7: blocky masks and id-equality masks
		uint32_t anyb = 0, crossed_cols = 0, amb = 0;
#pragma unroll 2
		for (int j = 0; j < 8; j++) {
			const int col = 32 * j + lane;
			const Col own = lds_col(raw + col * 16);
			const Col ny = lds_col(raw + ((col - 1) & 255) * 16);   // (x, y-1); unused garbage for y = 0
			const Col nx = lds_col(raw + ((col - 16) & 255) * 16);  // (x-1, y); unused garbage for x = 0
			uint32_t orf;
			const uint32_t bz = col_flags(flags, own, orf);
			smem_Bzp(s)[(2 * j + h + 1) * kBzStride + y + 1] = (uint16_t)bz;
			smem_Ey(s)[col] = (uint16_t)eq_mask16(own, ny);
			smem_Ex(s)[col] = (uint16_t)eq_mask16(own, nx);
			anyb |= bz;
			crossed_cols |= ((orf >> 1) & 1u) << j;
			amb |= orf;
		}
		amb = __reduce_or_sync(full, amb);
		uint32_t nq = 0, npl = 0;
		if (amb & kFlagBlocky) {  // else: no cube in the chunk (e.g. all air)
			uint32_t *di = p.desc_i + (size_t)c * kDescStrideI;
			__syncwarp();
			// B. run starts over z for the six passes
#pragma unroll 1
			for (int j = 0; j < 8; j++) col_starts(s, tab, raw, 2 * j + h, y, amb);
			__syncwarp();
			// C. masks over the merge axes; quads and non-empty words per group; scans in emission order
			transpose_blocky(s, lane);
			uint32_t tc[3], wofs[3], run = 0;
#pragma unroll
			for (int k = 0; k < 3; k++) {
				tc[k] = transpose_group(s, lane + 32 * k);
				const uint32_t inc = warp_incl_scan(tc[k], lane);  // quads (low half) and words (high half) at once
				wofs[k] = (run + inc - tc[k]) >> 16;
				run += __shfl_sync(full, inc, 31);
			}
			nq = run & 0xffffu;
			const uint32_t nwords = run >> 16;
			__syncwarp();  // Ey / Ex are dead: wl takes their place
#pragma unroll
			for (int k = 0; k < 3; k++)
				if (tc[k]) list_group_words(s, lane + 32 * k, wofs[k]);
			__syncwarp();
			// D. descriptors: lane <-> non-empty word (two rows), 32 words at a time
			uint32_t qbase = 0;
#pragma unroll 1
			for (uint32_t i0 = 0; i0 < nwords; i0 += 32) {
				const uint32_t i = i0 + (uint32_t)lane;
				uint32_t widx = 0, word = 0;
				if (i < nwords) {
					widx = smem_wl(s)[i];
					word = reinterpret_cast<const uint32_t *>(s.S)[widx];
				}
				const uint32_t cnt = (uint32_t)__popc(word);
				const uint32_t inc = warp_incl_scan(cnt, lane);
				if (word) emit_word_descs(s, widx, word, di + qbase + inc - cnt);
				qbase += __shfl_sync(full, inc, 31);
			}
		}
		if (amb & kFlagCrossed) {
			__syncwarp();  // wl is dead: Cz takes its place
			uint16_t *Cz = smem_Cz(s);
#pragma unroll 1
			for (int j = 0; j < 8; j++) {
				uint32_t cz = 0;
				if ((crossed_cols >> j) & 1u) cz = col_crossed(flags, lds_col(raw + (32 * j + lane) * 16));
				Cz[32 * j + lane] = (uint16_t)cz;
			}
			__syncwarp();
			const uint32_t mine = count_plants8(s, lane);
			const uint32_t inc = warp_incl_scan(mine, lane);
			npl = __shfl_sync(full, inc, 31);
			if (mine) emit_plant_descs(s, lane, p.desc_t + (size_t)c * kDescStrideT + (inc - mine));
		}
		if (lane == 0) {
			p.cnt_i[c] = nq;
			p.cnt_t[c] = npl * (uint32_t)kPlantQuads;
		}
	}
}

// ---- kernel 2: expansion -------------------------------------------------------------------
// Exclusive prefix of cnt[0..n) into pre[0..n], pre[n] = total; pre[k > n] = 0xffffffff so that a
// fixed-depth binary search never steps past n.  All kExpandThreads threads take part.
__device__ __forceinline__ void block_scan_counts(const uint32_t *cnt, uint32_t n, uint32_t *pre, uint32_t *warp_tot,
                                                  int tid) {
	const int lane = tid & 31, warp = tid >> 5;
	uint32_t v[kScanPerThread];
	const uint32_t first = (uint32_t)tid * kScanPerThread;
	uint32_t sum = 0;
#pragma unroll
	for (int k = 0; k < kScanPerThread; k += 4) {
		uint4 q = make_uint4(0, 0, 0, 0);
		if (first + k < n) q = *reinterpret_cast<const uint4 *>(cnt + first + k);  // the arrays are padded to kMaxLaunchChunks
		v[k] = first + k < n ? q.x : 0u;
		v[k + 1] = first + k + 1 < n ? q.y : 0u;
		v[k + 2] = first + k + 2 < n ? q.z : 0u;
		v[k + 3] = first + k + 3 < n ? q.w : 0u;
		sum += v[k] + v[k + 1] + v[k + 2] + v[k + 3];
	}
	const uint32_t inc = warp_incl_scan(sum, lane);
	if (lane == 31) warp_tot[warp] = inc;
	__syncthreads();
	uint32_t wt = lane < kExpandWarps ? warp_tot[lane] : 0u;
	const uint32_t winc = warp_incl_scan(wt, lane);
	const uint32_t wbase = __shfl_sync(0xffffffffu, winc - wt, warp);
	const uint32_t total = __shfl_sync(0xffffffffu, winc, kExpandWarps - 1);
	uint32_t run = wbase + inc - sum;
#pragma unroll
	for (int k = 0; k < kScanPerThread; k++) {
		const uint32_t i = first + k;
		pre[i] = i <= n ? run : 0xffffffffu;
		run += v[k];
	}
	if (tid < 4) pre[kMaxLaunchChunks + tid] = (kMaxLaunchChunks + tid) <= n ? total : 0xffffffffu;
	__syncthreads();
}

// largest c in [0, n) with pre[c] <= q  (q < pre[n]); `top` = highest power of two <= n
__device__ __forceinline__ uint32_t find_chunk(const uint32_t *pre, uint32_t q, uint32_t top) {
	uint32_t lo = 0;
	for (uint32_t step = top; step >= 1; step >>= 1)
		if (pre[lo + step] <= q) lo += step;
	return lo;
}

__global__ void __launch_bounds__(kExpandThreads, 1) expand_kernel(const ExpandParams p) {
	extern __shared__ __align__(128) unsigned char smem_raw_bytes[];
	TabEntry *tab = reinterpret_cast<TabEntry *>(smem_raw_bytes);
	uint32_t *pre_i = reinterpret_cast<uint32_t *>(smem_raw_bytes + 2048);
	uint32_t *pre_t = pre_i + kExpandPrefixWords;
	uint32_t *warp_tot = pre_t + kExpandPrefixWords;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	unsigned char *stage = smem_raw_bytes + 2048 + 2 * 4 * kExpandPrefixWords + 256 + (size_t)warp * (kStageBytes + 16);

	if (tid < 128) reinterpret_cast<uint4 *>(tab)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.table) + tid);
	pdl_wait();  // everything the analysis wrote is visible from here on
	block_scan_counts(p.cnt_i, p.n, pre_i, warp_tot, tid);
	block_scan_counts(p.cnt_t, p.n, pre_t, warp_tot, tid);
	const uint32_t tot_i = pre_i[p.n], tot_t = pre_t[p.n];
	u64 carry_i = 0, carry_t = 0;
	if (p.carry_in) {
		carry_i = p.carry_in[0];
		carry_t = p.carry_in[1];
	}
	// per-chunk ranges and the totals: spread over the grid
	for (uint32_t c = blockIdx.x * kExpandThreads + tid; c < p.n; c += gridDim.x * kExpandThreads) {
		Range ri, rt;
		ri.first_vertex = (carry_i + pre_i[c]) * 6ull; ri.vertex_count = (pre_i[c + 1] - pre_i[c]) * 6u; ri.reserved = 0;
		rt.first_vertex = (carry_t + pre_t[c]) * 6ull; rt.vertex_count = (pre_t[c + 1] - pre_t[c]) * 6u; rt.reserved = 0;
		p.ranges_i[c] = ri;
		p.ranges_t[c] = rt;
	}
	if (blockIdx.x == 0 && tid == 0) {
		p.totals[0] = carry_i + tot_i;
		p.totals[1] = carry_t + tot_t;
	}
	uint32_t top = 1;
	while (top * 2u <= p.n) top *= 2u;
	if (top == p.n) top >>= 1;  // the steps top, top/2, .., 1 then reach exactly [0, 2*top) >= [0, n)

	const bool fits_i = carry_i + tot_i <= p.cap_i, fits_t = carry_t + tot_t <= p.cap_t;
	const uint32_t tiles_i = fits_i ? (tot_i + kTileQuads - 1) / kTileQuads : 0u;
	const uint32_t tiles_t = fits_t ? (tot_t + kTileQuads - 1) / kTileQuads : 0u;
	const uint32_t gw = blockIdx.x * kExpandWarps + warp, nw = gridDim.x * kExpandWarps;
#pragma unroll 1
	for (uint32_t t = gw; t < tiles_i + tiles_t; t += nw) {
		const bool face = t < tiles_i;
		const uint32_t q0 = (face ? t : t - tiles_i) * kTileQuads;
		const uint32_t tot = face ? tot_i : tot_t;
		const uint32_t tn = min((uint32_t)kTileQuads, tot - q0);
		const uint32_t q = q0 + (uint32_t)lane;
		const bool act = (uint32_t)lane < tn;
		uint32_t d = 0, c = 0;
		long long ox = 0, oy = 0, oz = 0;
		if (act) {
			const uint32_t *pre = face ? pre_i : pre_t;
			c = find_chunk(pre, q, top);
			const uint32_t r = q - pre[c];
			d = face ? __ldg(p.desc_i + (size_t)c * kDescStrideI + r) : __ldg(p.desc_t + (size_t)c * kDescStrideT + (r >> 2));
			ox = __ldg(p.offs + 3 * (size_t)c);
			oy = __ldg(p.offs + 3 * (size_t)c + 1);
			oz = __ldg(p.offs + 3 * (size_t)c + 2);
		}
		unsigned char *g = (face ? p.out_i + (carry_i + q0) * (u64)kQuadBytes : p.out_t + (carry_t + q0) * (u64)kQuadBytes);
		unsigned char *sp = stage + ((uintptr_t)g & 8u);
		if (lane == 0) bulk_wait_read_all();  // the staging tile is free once the last bulk store has read it
		__syncwarp();
		if (act) {
			if (face) expand_face_desc(tab, ox, oy, oz, d, sp + lane * kQuadBytes);
			else expand_plant_desc(tab, ox, oy, oz, d, (int)(q & 3u), sp + lane * kQuadBytes);
		}
		fence_proxy_async_smem();
		__syncwarp();
		if (lane == 0) store_tile(g, sp, tn * kQuadBytes);
	}
	if (lane == 0) bulk_wait_all();  // all bulk stores of this warp have landed before it exits
}

}  // namespace mm
