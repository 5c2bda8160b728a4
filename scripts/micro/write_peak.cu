// write_peak.cu -- how fast can this GPU write a linear stream?  (context for the mesher's roofline: 91 % of its
// traffic is writes, while the roofline's denominator, MEASURED_PEAKS.json hbm_gbs, is a copy = half reads.)
// Persistent CTAs, every warp stores tiles of 6912 bytes (the mesher's tile) from shared memory with cp.async.bulk,
// the whole grid sweeping the buffer front to back like expand_kernel does; no computation at all.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o write_peak write_peak.cu && ./write_peak
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

constexpr int kTile = 6912;

__device__ __forceinline__ void bulk_store(void *g, const void *s, uint32_t bytes) {
	const uint32_t sa = (uint32_t)__cvta_generic_to_shared(s);
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(sa), "r"(bytes) : "memory");
	asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

template <int WARPS, int DEPTH>
__global__ void __launch_bounds__(WARPS * 32) sweep(unsigned char *out, unsigned long long tiles) {
	extern __shared__ __align__(128) unsigned char sm[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	unsigned char *tile = sm + (size_t)warp * kTile;
	for (int i = lane; i < kTile / 4; i += 32) reinterpret_cast<uint32_t *>(tile)[i] = 0x01010101u * (uint32_t)(warp + 1);
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	__syncwarp();
	const unsigned long long gw = (unsigned long long)blockIdx.x * WARPS + warp, nw = (unsigned long long)gridDim.x * WARPS;
	if (lane == 0) {
		for (unsigned long long t = gw; t < tiles; t += nw) {
			// at most DEPTH stores in flight per warp (the mesher: 1, it refills the tile in between)
			if (DEPTH == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
			else if (DEPTH == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
			else asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
			bulk_store(out + t * kTile, tile, kTile);
		}
		asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
	}
}

template <int WARPS, int DEPTH>
static void run(unsigned char *buf, size_t bytes, int sms, int ctas_per_sm) {
	const unsigned long long tiles = bytes / kTile;
	const size_t smem = (size_t)WARPS * kTile;
	cudaFuncSetAttribute(sweep<WARPS, DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e9f;
	for (int r = 0; r < 6; r++) {
		cudaEventRecord(e0);
		sweep<WARPS, DEPTH><<<sms * ctas_per_sm, WARPS * 32, smem>>>(buf, tiles);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (r && ms < best) best = ms;
	}
	printf("{\"warps_per_cta\": %d, \"ctas_per_sm\": %d, \"stores_in_flight_per_warp\": %d, \"write_gbs\": %.1f, \"error\": \"%s\"}\n", WARPS, ctas_per_sm, DEPTH,
	       (double)tiles * kTile / (best * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
	int sms = 0;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
	const size_t bytes = (size_t)8 << 30;  // 8 GiB: far beyond the 126 MB L2
	unsigned char *buf = nullptr;
	if (cudaMalloc(&buf, bytes) != cudaSuccess) return 1;
	run<12, 1>(buf, bytes, sms, 1);
	run<12, 2>(buf, bytes, sms, 1);
	run<12, 4>(buf, bytes, sms, 1);
	run<8, 4>(buf, bytes, sms, 1);
	run<16, 4>(buf, bytes, sms, 1);
	run<4, 4>(buf, bytes, sms, 2);
	run<8, 4>(buf, bytes, sms, 2);
	cudaFree(buf);
	return 0;
}
