// mesh_kernel.cuh -- the sm_100a batch mesher kernel, v2 (device only).
//
// Persistent CTAs of 8 warps; every WARP is an independent chunk worker that draws chunk
// tickets in order and runs the phases of mesh_core.cuh with its 32 lanes:
//   A. 8 x (one 16-byte streaming load per lane) = the 4 KiB chunk; blocky masks, id-equality
//      masks (neighbour columns arrive by warp shuffle)
//   B. visibility / run starts of the six passes as bit operations on 16-bit z-masks
//   C. in-register 16x16 bit transposes -> row masks in emission order, popc -> counts,
//      warp shuffle scan -> quad offsets inside the chunk
//   D. chunk base in the packed streams: every chunk publishes its counts in a status word;
//      groups of 32 chunks and of 1024 chunks are summed by whichever warp arrives last
//      (one atomic per chunk on a per-group counter), so a chunk's exclusive prefix is
//      3 coalesced 32-lane loads (chunks of its group, groups of its super-group, super-groups)
//      -- single pass, tightly packed, chunk order, no serial chain between chunks
//   E. lane <-> quad: 216 B per lane into the warp's staging tile, then one cp.async.bulk
//      (TMA) shared -> global store per tile, overlapped with the search for the next tile
// No CTA-wide barrier after set-up; no tensor cores: the path is a scan/compaction bound by
// the vertex-stream writes.
#pragma once
#include <cuda_runtime.h>
#include "mesh_core.cuh"

namespace mm {

constexpr int kWarpsPerCta = 8;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr uint32_t kMaxLaunchChunks = 32768;  // 32 super-groups of 32 groups of 32 chunks
constexpr size_t kCtaTableBytes = 2048 + 256;
constexpr size_t kCtaSmemBytes = kCtaTableBytes + kWarpsPerCta * sizeof(WarpSmem);

struct KernelParams {
	const unsigned char *chunks;  // n * 4096
	const long long *offs;        // n * 3
	uint32_t n;                   // <= kMaxLaunchChunks
	uint32_t epoch;               // 1..16383, tags the status words of this launch
	const TabEntry *table;        // 256 entries
	const unsigned char *flags;   // 256 flag bytes (kFlagBlocky / kFlagCrossed)
	unsigned char *out_i;         // intransparent vertex stream
	unsigned char *out_t;         // transparent vertex stream
	u64 cap_i, cap_t;             // arena capacities in quads (216 B units)
	Range *ranges_i, *ranges_t;   // n entries each
	u64 *st_chunk;                // status words: per chunk, per group of 32, per super-group of 1024
	u64 *st_group;
	u64 *st_super;
	uint32_t *cnt_group;          // arrival counters (zero between launches)
	uint32_t *cnt_super;
	u64 *ticket;                  // monotonically increasing across launches
	u64 ticket_base;              // its value at launch
	const u64 *carry_in;          // quads already in the streams (earlier launches of the batch) or NULL
	u64 *totals;                  // [0] quads in out_i, [1] quad slots in out_t after this launch
};

// status word: epoch(14) | t quad slots(25) | i quads(25)
constexpr u64 kCntMask = (1ull << 25) - 1;
__device__ __forceinline__ u64 pack_status(uint32_t epoch, uint32_t i, uint32_t t) {
	return ((u64)epoch << 50) | ((u64)t << 25) | (u64)i;
}
__device__ __forceinline__ u64 ld_status(const u64 *p) {
	u64 v;
	asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_status(u64 *p, u64 v) {
	asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t arrive(uint32_t *p) {  // release our status store, acquire the others'
	uint32_t old;
	asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(p) : "memory");
	return old;
}
__device__ __forceinline__ u64 wait_status(const u64 *p, uint32_t epoch) {
	u64 w = ld_status(p);
	while ((uint32_t)(w >> 50) != epoch) {
		__nanosleep(64);
		w = ld_status(p);
	}
	return w;
}
__device__ __forceinline__ Col ld_col(const unsigned char *p) {  // read-once input: keep it out of L1
	Col c;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	             : "=r"(c.w[0]), "=r"(c.w[1]), "=r"(c.w[2]), "=r"(c.w[3])
	             : "l"(p));
	return c;
}
__device__ __forceinline__ uint32_t warp_sum(uint32_t v) { return __reduce_add_sync(0xffffffffu, v); }

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

// ---- phase D: publish this chunk's counts, obtain its exclusive prefix --------------------
__device__ __forceinline__ void sum_status(u64 w, uint32_t &si, uint32_t &st) {
	si = warp_sum((uint32_t)(w & kCntMask));
	st = warp_sum((uint32_t)((w >> 25) & kCntMask));
}

__device__ __forceinline__ void publish_and_prefix(const KernelParams &p, uint32_t c, uint32_t nq, uint32_t nts,
                                                   int lane, u64 &base_i, u64 &base_t) {
	const uint32_t g = c >> 5, r = c & 31u, sg = g >> 5, gs = g & 31u;
	uint32_t old = 0;
	if (lane == 0) {
		st_status(p.st_chunk + c, pack_status(p.epoch, nq, nts));
		old = arrive(p.cnt_group + g);
	}
	old = __shfl_sync(0xffffffffu, old, 0);
	const uint32_t size1 = min(32u, p.n - (g << 5));
	if (old == size1 - 1u) {  // last chunk of its group to arrive: publish the group sum
		const u64 w = (uint32_t)lane < size1 ? wait_status(p.st_chunk + (g << 5) + lane, p.epoch) : 0ull;
		uint32_t si, st;
		sum_status(w, si, st);
		uint32_t old2 = 0;
		if (lane == 0) {
			st_status(p.st_group + g, pack_status(p.epoch, si, st));
			p.cnt_group[g] = 0;  // ready for the next launch
			old2 = arrive(p.cnt_super + sg);
		}
		old2 = __shfl_sync(0xffffffffu, old2, 0);
		const uint32_t size2 = min(32u, ((p.n + 31u) >> 5) - (sg << 5));
		if (old2 == size2 - 1u) {  // last group of its super-group
			const u64 w2 = (uint32_t)lane < size2 ? wait_status(p.st_group + (sg << 5) + lane, p.epoch) : 0ull;
			sum_status(w2, si, st);
			if (lane == 0) {
				st_status(p.st_super + sg, pack_status(p.epoch, si, st));
				p.cnt_super[sg] = 0;
			}
		}
	}
	// exclusive prefix = earlier chunks of this group + earlier groups of this super-group
	// + earlier super-groups (each at most 31 words, one per lane)
	uint32_t ai = 0, at = 0;
	if ((uint32_t)lane < r) {
		const u64 w = wait_status(p.st_chunk + (g << 5) + lane, p.epoch);
		ai += (uint32_t)(w & kCntMask);
		at += (uint32_t)((w >> 25) & kCntMask);
	}
	if ((uint32_t)lane < gs) {
		const u64 w = wait_status(p.st_group + (sg << 5) + lane, p.epoch);
		ai += (uint32_t)(w & kCntMask);
		at += (uint32_t)((w >> 25) & kCntMask);
	}
	if ((uint32_t)lane < sg) {
		const u64 w = wait_status(p.st_super + lane, p.epoch);
		ai += (uint32_t)(w & kCntMask);
		at += (uint32_t)((w >> 25) & kCntMask);
	}
	u64 ci = 0, ct = 0;
	if (p.carry_in) {
		ci = p.carry_in[0];
		ct = p.carry_in[1];
	}
	base_i = ci + warp_sum(ai);
	base_t = ct + warp_sum(at);
}

// ---- the kernel -------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 2) mesh_batch_kernel(const KernelParams p) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	TabEntry *tab = reinterpret_cast<TabEntry *>(smem_raw);
	unsigned char *flags = smem_raw + 2048;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	WarpSmem &s = *reinterpret_cast<WarpSmem *>(smem_raw + kCtaTableBytes + (size_t)warp * sizeof(WarpSmem));
	const unsigned full = 0xffffffffu;

	// block table + flag bytes -> shared memory, once per CTA
	if (tid < 128) reinterpret_cast<uint4 *>(tab)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.table) + tid);
	if (tid < 16) reinterpret_cast<uint4 *>(flags)[tid] = __ldg(reinterpret_cast<const uint4 *>(p.flags) + tid);
	init_warp_smem(s, lane);
	__syncthreads();

	const int h = lane >> 4, y = lane & 15;
	for (;;) {
		// the staging tile (= Ey/Ex scratch) is free once the last bulk store has read it
		if (lane == 0) bulk_wait_read_all();
		u64 ticket = 0;
		if (lane == 0) ticket = atomicAdd(p.ticket, 1ull) - p.ticket_base;
		ticket = __shfl_sync(full, ticket, 0);  // also orders the previous chunk's smem reads before this one's writes
		if (ticket >= (u64)p.n) break;
		const uint32_t c = (uint32_t)ticket;
		const unsigned char *chunk = p.chunks + (size_t)c * 4096u;

		// A. the chunk: lane (h, y) owns columns (x = 2j + h, y), j = 0..7
		Col cols[8];
#pragma unroll
		for (int j = 0; j < 8; j++) cols[j] = ld_col(chunk + (size_t)(32 * j + lane) * 16u);
		phase_pos(s, lane, p.offs + 3 * (size_t)c);
		uint32_t anyb = 0, crossed_cols = 0;
#pragma unroll
		for (int j = 0; j < 8; j++) {
			Col ny, nx;  // columns (x, y-1) and (x-1, y)
#pragma unroll
			for (int k = 0; k < 4; k++) {
				ny.w[k] = __shfl_up_sync(full, cols[j].w[k], 1);
				const uint32_t send = h ? (j > 0 ? cols[j > 0 ? j - 1 : 0].w[k] : 0u) : cols[j].w[k];
				nx.w[k] = __shfl_xor_sync(full, send, 16);
			}
			uint32_t sum;
			const uint32_t bz = col_flags(flags, cols[j], sum);
			const int col = 32 * j + lane;
			s.Bzp[(2 * j + h + 1) * kBzStride + y + 1] = (uint16_t)bz;
			smem_Ey(s)[col] = (uint16_t)eq_mask16(cols[j], ny);
			smem_Ex(s)[col] = (uint16_t)eq_mask16(cols[j], nx);
			anyb |= bz;
			crossed_cols |= ((sum >> 5) ? 1u : 0u) << j;
		}
		uint32_t nq = 0, npl = 0;
		if (__any_sync(full, (anyb | crossed_cols) != 0u)) {  // else: nothing drawn (e.g. an all-air chunk)
			__syncwarp();
			// B. run starts over z; crossed masks
#pragma unroll
			for (int j = 0; j < 8; j++) col_starts(s, tab, chunk, 2 * j + h, y);
			if (__any_sync(full, crossed_cols != 0u)) {
				uint32_t mine = 0;
#pragma unroll
				for (int j = 0; j < 8; j++) {
					const uint32_t cz = ((crossed_cols >> j) & 1u) ? col_crossed(flags, cols[j]) : 0u;
					s.Cz[32 * j + lane] = (uint16_t)cz;
					mine += (uint32_t)__popc(cz);
				}
				npl = warp_sum(mine);
			}
			__syncwarp();
			// C. masks over the merge axes; per-group quad counts; scan in emission order
			transpose_blocky(s, lane);
			uint32_t run = 0;
#pragma unroll
			for (int k = 0; k < 3; k++) {
				const uint32_t cnt = transpose_group(s, lane + 32 * k);
				uint32_t inc = cnt;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const uint32_t o = __shfl_up_sync(full, inc, d);
					if (lane >= d) inc += o;
				}
				s.gpre[lane + 32 * k] = (uint16_t)(run + inc - cnt);
				run += __shfl_sync(full, inc, 31);
			}
			nq = run;
			__syncwarp();
		}
		const uint32_t nts = npl * (uint32_t)kPlantQuads;

		// D. where this chunk's vertices go
		u64 base_i, base_t;
		publish_and_prefix(p, c, nq, nts, lane, base_i, base_t);
		if (lane == 0) {
			Range ri, rt;
			ri.first_vertex = base_i * 6ull; ri.vertex_count = nq * 6u; ri.reserved = 0;
			rt.first_vertex = base_t * 6ull; rt.vertex_count = nts * 6u; rt.reserved = 0;
			p.ranges_i[c] = ri;
			p.ranges_t[c] = rt;
			if (c == p.n - 1) {
				p.totals[0] = base_i + nq;
				p.totals[1] = base_t + nts;
			}
		}

		// E1. intransparent stream: the six face passes
		if (nq != 0 && base_i + nq <= p.cap_i) {
			for (uint32_t t0 = 0; t0 < nq; t0 += kTileQuads) {
				const uint32_t tn = min((uint32_t)kTileQuads, nq - t0);
				int m = 0, c2 = 0, st = 0;
				if ((uint32_t)lane < tn) find_quad(s, t0 + lane, m, c2, st);
				if (lane == 0) bulk_wait_read_all();  // staging tile free again
				__syncwarp();
				unsigned char *g = p.out_i + (base_i + t0) * (u64)kQuadBytes;
				unsigned char *sp = s.stage + ((uintptr_t)g & 8u);
				if ((uint32_t)lane < tn) expand_face_quad(s, tab, chunk, m, c2, st, sp + lane * kQuadBytes);
				fence_proxy_async_smem();
				__syncwarp();
				if (lane == 0) store_tile(g, sp, tn * kQuadBytes);
			}
		}
		// E2. transparent stream: crossed plants, 4 quad slots each
		if (npl != 0 && base_t + nts <= p.cap_t) {
			__syncwarp();  // S is no longer read: reuse it for the plant prefix
			uint16_t *cpre = smem_cpre(s);
			uint32_t cnt[8], mine = 0;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				cnt[k] = (uint32_t)__popc((uint32_t)s.Cz[lane * 8 + k]);
				mine += cnt[k];
			}
			uint32_t inc = mine;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t o = __shfl_up_sync(full, inc, d);
				if (lane >= d) inc += o;
			}
			uint32_t ex = inc - mine;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				cpre[lane * 8 + k] = (uint16_t)ex;
				ex += cnt[k];
			}
			__syncwarp();
			for (uint32_t t0 = 0; t0 < nts; t0 += kTileQuads) {
				const uint32_t tn = min((uint32_t)kTileQuads, nts - t0);
				int cell = 0;
				if ((uint32_t)lane < tn) cell = find_plant(s, (t0 + lane) >> 2);
				if (lane == 0) bulk_wait_read_all();
				__syncwarp();
				unsigned char *g = p.out_t + (base_t + t0) * (u64)kQuadBytes;
				unsigned char *sp = s.stage + ((uintptr_t)g & 8u);
				if ((uint32_t)lane < tn) expand_plant_quad(s, tab, chunk, cell, (int)((t0 + lane) & 3u), sp + lane * kQuadBytes);
				fence_proxy_async_smem();
				__syncwarp();
				if (lane == 0) store_tile(g, sp, tn * kQuadBytes);
			}
		}
	}
	if (lane == 0) bulk_wait_all();  // all bulk stores of this warp have landed before it exits
}

}  // namespace mm
