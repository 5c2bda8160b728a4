// mesh_kernel.cuh -- the sm_100a batch mesher kernel (device only).
//
// One persistent CTA (128 threads) takes chunk tickets in order.  Per chunk:
//   1. 4 KiB of block ids -> shared memory with two 16-byte read-only loads per thread
//   2. mesh_core.cuh phases: row masks, visibility, run starts, counts
//   3. CTA-wide exclusive scan of the 6 x 256 row counts in emission order
//   4. chunk base = exclusive prefix over all earlier chunks, obtained in the same
//      launch by decoupled look-back over a per-chunk status word (single pass, the
//      streams come out tightly packed in chunk order)
//   5. descriptors -> one thread per quad writes 216 B into a staging tile ->
//      cp.async.bulk (TMA) shared -> global store of the tile, overlapped with the
//      next tile / next chunk
// No tensor cores: the path is a scan/compaction, bound by the vertex-stream writes.
#pragma once
#include <cuda_runtime.h>
#include "mesh_core.cuh"

namespace mm {

struct KernelParams {
	const unsigned char *chunks;  // n * 4096
	const long long *offs;        // n * 3
	uint32_t n;
	const TabEntry *table;        // 256 entries
	unsigned char *out_i;         // intransparent vertex stream
	unsigned char *out_t;         // transparent vertex stream
	u64 cap_i, cap_t;             // arena capacities in quads (216 B units)
	Range *ranges_i, *ranges_t;
	u64 *status;                  // n look-back words for this launch (all zero at launch)
	u64 *status_next;             // the other buffer: slot c is zeroed for the next launch
	u64 *ticket;                  // monotonically increasing across launches
	u64 ticket_base;              // its value at launch
	u64 *totals;                  // [0] quads in out_i, [1] quad slots in out_t
};

// look-back word: flag(2) | t quad slots (31) | i quads (31); counts saturate at kSat
constexpr u64 kFlagAgg = 1ull << 62, kFlagPre = 2ull << 62;
constexpr u64 kSat = 0x7fffffffull;

__device__ __forceinline__ u64 ld_status(const u64 *p) {
	u64 v;
	asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_status(u64 *p, u64 v) {
	asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 pack_counts(u64 i, u64 t) {
	return (i > kSat ? kSat : i) | ((t > kSat ? kSat : t) << 31);
}
__device__ __forceinline__ uint4 ld_stream16(const uint4 *p) {  // read-once input: keep it out of L1
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	             : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
	             : "l"(p));
	return r;
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

// Called by ONE thread after the tile is complete and fenced.  `g` is the byte address
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

// ---- CTA-wide exclusive scan over u16 counts -----------------------------------------
// a[0 .. kThreads*PER) counts -> exclusive prefix in place; returns the total.
template <int PER>
__device__ __forceinline__ uint32_t block_exclusive_scan_u16(uint16_t *a, Smem &s, int tid) {
	uint32_t v[PER];
	uint32_t sum = 0;
#pragma unroll
	for (int i = 0; i < PER; i += 2) {
		const uint32_t w = *reinterpret_cast<const uint32_t *>(a + tid * PER + i);
		v[i] = w & 0xffffu;
		v[i + 1] = w >> 16;
		sum += v[i] + v[i + 1];
	}
	const int lane = tid & 31, warp = tid >> 5;
	uint32_t inc = sum;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
		if (lane >= d) inc += o;
	}
	if (lane == 31) s.warp_sums[warp] = inc;
	__syncthreads();
	uint32_t base = 0, total = 0;
#pragma unroll
	for (int w = 0; w < kThreads / 32; w++) {
		const uint32_t ws = s.warp_sums[w];
		if (w < warp) base += ws;
		total += ws;
	}
	uint32_t run = base + inc - sum;
#pragma unroll
	for (int i = 0; i < PER; i += 2) {
		const uint32_t lo = run;
		run += v[i];
		const uint32_t hi = run;
		run += v[i + 1];
		*reinterpret_cast<uint32_t *>(a + tid * PER + i) = lo | (hi << 16);
	}
	__syncthreads();
	return total;
}

// ---- decoupled look-back (warp 0) ----------------------------------------------------
__device__ __forceinline__ void lookback_and_publish(const KernelParams &p, Smem &s, uint32_t c,
                                                     uint32_t nq, uint32_t nts, int lane) {
	u64 ex_i = 0, ex_t = 0;
	if (c == 0) {
		if (lane == 0) st_status(p.status, kFlagPre | pack_counts(nq, nts));
	} else {
		if (lane == 0) st_status(p.status + c, kFlagAgg | pack_counts(nq, nts));
		long long j = (long long)c - 1;
		for (;;) {
			const long long idx = j - lane;
			u64 w = kFlagPre;  // virtual chunk -1: prefix 0
			if (idx >= 0) {
				w = ld_status(p.status + idx);
				while ((w >> 62) == 0) {
					__nanosleep(32);
					w = ld_status(p.status + idx);
				}
			}
			const unsigned pm = __ballot_sync(0xffffffffu, (w >> 62) == 2);
			const int first = pm ? (__ffs((int)pm) - 1) : 32;
			u64 vi = (lane <= first) ? (w & kSat) : 0;
			u64 vt = (lane <= first) ? ((w >> 31) & kSat) : 0;
#pragma unroll
			for (int d = 16; d >= 1; d >>= 1) {
				vi += __shfl_xor_sync(0xffffffffu, vi, d);
				vt += __shfl_xor_sync(0xffffffffu, vt, d);
			}
			ex_i += vi;
			ex_t += vt;
			if (pm) break;
			j -= 32;
		}
		if (lane == 0) st_status(p.status + c, kFlagPre | pack_counts(ex_i + nq, ex_t + nts));
	}
	if (lane == 0) {
		p.status_next[c] = 0;
		s.base_q = ex_i;
		s.base_t = ex_t;
		Range ri, rt;
		ri.first_vertex = ex_i * 6ull; ri.vertex_count = nq * 6u; ri.reserved = 0;
		rt.first_vertex = ex_t * 6ull; rt.vertex_count = nts * 6u; rt.reserved = 0;
		p.ranges_i[c] = ri;
		p.ranges_t[c] = rt;
		if (c == p.n - 1) {
			p.totals[0] = ex_i + nq;
			p.totals[1] = ex_t + nts;
		}
	}
}

// ---- the kernel -------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 4) mesh_batch_kernel(const KernelParams p) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	Smem &s = *reinterpret_cast<Smem *>(smem_raw);
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

	// block table -> shared memory, once per CTA (2 KiB)
	{
		const uint4 *t4 = reinterpret_cast<const uint4 *>(p.table);
		reinterpret_cast<uint4 *>(s.tab)[tid] = __ldg(t4 + tid);
	}

	for (;;) {
		__syncthreads();  // previous chunk fully consumed (desc, masks, ticket)
		if (tid == 0) s.ticket = atomicAdd(p.ticket, 1ull) - p.ticket_base;
		__syncthreads();
		const u64 ticket = s.ticket;
		if (ticket >= (u64)p.n) break;
		const uint32_t c = (uint32_t)ticket;

		// 1. stage the chunk
		const uint4 *src = reinterpret_cast<const uint4 *>(p.chunks + (size_t)c * 4096u);
		const uint4 v0 = ld_stream16(src + tid), v1 = ld_stream16(src + tid + kThreads);
		reinterpret_cast<uint4 *>(s.raw)[tid] = v0;
		reinterpret_cast<uint4 *>(s.raw)[tid + kThreads] = v1;
		phase_pos(s, tid, p.offs + 3 * (size_t)c);
		if (tid == 0) s.any_crossed = 0;
		const uint32_t w = v0.x;
		const bool local_uniform = (v0.y == w) & (v0.z == w) & (v0.w == w) & (v1.x == w) & (v1.y == w) &
		                           (v1.z == w) & (v1.w == w) & (w == (w & 0xffu) * 0x01010101u);
		__syncthreads();
		const bool uniform =
		    __syncthreads_and(local_uniform && w == reinterpret_cast<const uint32_t *>(s.raw)[0]);

		// 2./3. masks, counts, scan
		uint32_t nq = 0, npl = 0;
		const bool trivial = uniform && (s.tab[w & 0xffu].y >> 16) == 0u;  // e.g. an all-air chunk
		if (!trivial) {
			uint32_t cr = 0;
			const RowRegs g0 = phase_rows(s, tid, cr);
			const RowRegs g1 = phase_rows(s, tid + kThreads, cr);
			if (cr) s.any_crossed = 1;
			__syncthreads();
			phase_vis(s, tid, g0);
			phase_vis(s, tid + kThreads, g1);
			const bool anyc = s.any_crossed != 0;
			if (anyc) {
				phase_crossed_cols(s, tid);
				phase_crossed_cols(s, tid + kThreads);
			}
			__syncthreads();
			nq = block_exclusive_scan_u16<12>(s.pre, s, tid);
			if (anyc) npl = block_exclusive_scan_u16<2>(s.cpre, s, tid);
		}
		const uint32_t nts = npl * (uint32_t)kPlantQuads;

		// 4. chunk base offsets
		if (warp == 0) lookback_and_publish(p, s, c, nq, nts, lane);
		__syncthreads();
		const u64 base_q = s.base_q, base_t = s.base_t;

		// 5a. intransparent stream: the six face passes
		if (nq != 0 && base_q + nq <= p.cap_i) {
			for (uint32_t w0 = 0; w0 < nq; w0 += kDescCap) {
				phase_desc_row(s, tid, w0);
				phase_desc_row(s, tid + kThreads, w0);
				__syncthreads();
				const uint32_t wend = min(nq, w0 + (uint32_t)kDescCap);
				for (uint32_t t0 = w0; t0 < wend; t0 += kTileQuads) {
					const uint32_t tn = min((uint32_t)kTileQuads, wend - t0);
					if (tid == 0) bulk_wait_read_all();  // staging tile free again
					__syncthreads();
					unsigned char *g = p.out_i + (base_q + t0) * (u64)kQuadBytes;
					unsigned char *sp = s.stage + ((uintptr_t)g & 8u);
					if ((uint32_t)tid < tn) expand_face_quad(s, s.desc[t0 - w0 + tid], sp + tid * kQuadBytes);
					fence_proxy_async_smem();
					__syncthreads();
					if (tid == 0) store_tile(g, sp, tn * kQuadBytes);
				}
			}
		}
		// 5b. transparent stream: crossed plants, 4 quad slots each
		if (npl != 0 && base_t + nts <= p.cap_t) {
			for (uint32_t w0 = 0; w0 < npl; w0 += kDescCap) {
				__syncthreads();  // descriptors of the previous round / stream are consumed
				phase_desc_col(s, tid, w0);
				phase_desc_col(s, tid + kThreads, w0);
				__syncthreads();
				const uint32_t qend = min(npl, w0 + (uint32_t)kDescCap) * kPlantQuads;
				for (uint32_t t0 = w0 * kPlantQuads; t0 < qend; t0 += kTileQuads) {
					const uint32_t tn = min((uint32_t)kTileQuads, qend - t0);
					if (tid == 0) bulk_wait_read_all();
					__syncthreads();
					unsigned char *g = p.out_t + (base_t + t0) * (u64)kQuadBytes;
					unsigned char *sp = s.stage + ((uintptr_t)g & 8u);
					if ((uint32_t)tid < tn) {
						const uint32_t q = t0 + tid;
						expand_plant_quad(s, s.desc[(q >> 2) - w0], (int)(q & 3u), sp + tid * kQuadBytes);
					}
					fence_proxy_async_smem();
					__syncthreads();
					if (tid == 0) store_tile(g, sp, tn * kQuadBytes);
				}
			}
		}
	}
	if (tid == 0) bulk_wait_all();  // all bulk stores of this CTA have landed before it exits
}

}  // namespace mm
