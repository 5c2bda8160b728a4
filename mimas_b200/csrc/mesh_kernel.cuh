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
// Arrival counter of a group.  Relaxed on purpose: the counter only elects the warp that sums the
// group; every status word it then reads carries its own epoch tag and is waited for individually.
__device__ __forceinline__ uint32_t arrive(uint32_t *p) {
	uint32_t old;
	asm volatile("atom.relaxed.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(p) : "memory");
	return old;
}
__device__ __forceinline__ u64 wait_status(const u64 *p, uint32_t epoch, uint32_t max_ns) {
	u64 w = ld_status(p);
	uint32_t ns = 32;
	while ((uint32_t)(w >> 50) != epoch) {
		__nanosleep(ns);
		if (ns < max_ns) ns <<= 1;
		w = ld_status(p);
	}
	return w;
}
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
	const uint32_t max_ns = (nq | nts) ? 256u : 2048u;  // an empty chunk is in no hurry
	const uint32_t g = c >> 5, r = c & 31u, sg = g >> 5, gs = g & 31u;
	uint32_t old = 0;
	if (lane == 0) {
		st_status(p.st_chunk + c, pack_status(p.epoch, nq, nts));
		old = arrive(p.cnt_group + g);
	}
	old = __shfl_sync(0xffffffffu, old, 0);
	const uint32_t size1 = min(32u, p.n - (g << 5));
	if (old == size1 - 1u) {  // last chunk of its group to arrive: publish the group sum
		const u64 w = (uint32_t)lane < size1 ? wait_status(p.st_chunk + (g << 5) + lane, p.epoch, 64u) : 0ull;
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
			const u64 w2 = (uint32_t)lane < size2 ? wait_status(p.st_group + (sg << 5) + lane, p.epoch, 64u) : 0ull;
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
		const u64 w = wait_status(p.st_chunk + (g << 5) + lane, p.epoch, max_ns);
		ai += (uint32_t)(w & kCntMask);
		at += (uint32_t)((w >> 25) & kCntMask);
	}
	if ((uint32_t)lane < gs) {
		const u64 w = wait_status(p.st_group + (sg << 5) + lane, p.epoch, max_ns);
		ai += (uint32_t)(w & kCntMask);
		at += (uint32_t)((w >> 25) & kCntMask);
	}
	if ((uint32_t)lane < sg) {
		const u64 w = wait_status(p.st_super + lane, p.epoch, max_ns);
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
	extern __shared__ __align__(128) unsigned char smem_raw_bytes[];
	TabEntry *tab = reinterpret_cast<TabEntry *>(smem_raw_bytes);
	unsigned char *flags = smem_raw_bytes + 2048;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	WarpSmem &s = *reinterpret_cast<WarpSmem *>(smem_raw_bytes + kCtaTableBytes + (size_t)warp * sizeof(WarpSmem));
	const unsigned full = 0xffffffffu;

	// block table + flag bytes -> shared memory, once per CTA
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
	for (;;) {
		u64 ticket = 0;
		if (lane == 0) {
			bulk_wait_read_all();  // the staging tile (= raw chunk + Ey/Ex) is free once the last bulk store has read it
			ticket = atomicAdd(p.ticket, 1ull) - p.ticket_base;
		}
		ticket = __shfl_sync(full, ticket, 0);  // also orders the previous chunk's smem reads before this one's writes
		if (ticket >= (u64)p.n) break;
		const uint32_t c = (uint32_t)ticket;
		const unsigned char *chunk = p.chunks + (size_t)c * 4096u;
		if (lane == 0) bulk_load_g2s(raw, chunk, 4096u, &s.mbar);  // TMA: 4 KiB chunk -> shared memory
		phase_pos(s, lane, p.offs + 3 * (size_t)c);
		zero_border(s, lane);
		mbar_wait(&s.mbar, parity);
		parity ^= 1u;

		// A. lane (h, y) owns columns (x = 2j + h, y), j = 0..7: blocky masks and id-equality masks
		uint32_t anyb = 0, crossed_cols = 0, ncrossed = 0;
#pragma unroll 2
		for (int j = 0; j < 8; j++) {
			const int col = 32 * j + lane;
			const Col own = lds_col(raw + col * 16);
			const Col ny = lds_col(raw + (col - 1) * 16);   // (x, y-1); unused garbage for y = 0
			const Col nx = lds_col(raw + (col - 16) * 16);  // (x-1, y); unused garbage for x = 0
			uint32_t sum;
			const uint32_t bz = col_flags(flags, own, sum);
			smem_Bzp(s)[(2 * j + h + 1) * kBzStride + y + 1] = (uint16_t)bz;
			smem_Ey(s)[col] = (uint16_t)eq_mask16(own, ny);
			smem_Ex(s)[col] = (uint16_t)eq_mask16(own, nx);
			anyb |= bz;
			crossed_cols |= ((sum >> 5) ? 1u : 0u) << j;
			ncrossed += sum >> 5;
		}
		uint32_t nq = 0, npl = 0;
		if (__any_sync(full, (anyb | crossed_cols) != 0u)) {  // else: nothing drawn (e.g. an all-air chunk)
			npl = warp_sum(ncrossed);
			__syncwarp();
			// B. run starts over z for the six passes
#pragma unroll 1
			for (int j = 0; j < 8; j++) col_starts(s, tab, raw, 2 * j + h, y);
			__syncwarp();
			// C. masks over the merge axes; per-group quad counts; scan in emission order
			transpose_blocky(s, lane);
			__syncwarp();  // Bzp is dead from here on: its bytes become rowpre
			uint32_t run = 0;
#pragma unroll 1
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
#pragma unroll 1
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
			__syncwarp();  // S is no longer read: reuse it for the crossed masks and their prefix
			uint16_t *Cz = smem_Cz(s), *cpre = smem_cpre(s);
#pragma unroll 1
			for (int j = 0; j < 8; j++) {
				uint32_t cz = 0;
				if ((crossed_cols >> j) & 1u) {
					const uint4 v = __ldg(reinterpret_cast<const uint4 *>(chunk) + 32 * j + lane);
					Col cc;
					cc.w[0] = v.x; cc.w[1] = v.y; cc.w[2] = v.z; cc.w[3] = v.w;
					cz = col_crossed(flags, cc);
				}
				Cz[32 * j + lane] = (uint16_t)cz;
			}
			__syncwarp();
			uint32_t cnt[8], mine = 0;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				cnt[k] = (uint32_t)__popc((uint32_t)Cz[lane * 8 + k]);
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
#pragma unroll 1
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
