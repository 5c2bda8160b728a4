// TEST INFRASTRUCTURE -- host emulation of the mesher kernel's control flow.
//
// Compiles mimas_b200/csrc/mesh_core.cuh (the per-lane phase functions the CUDA kernel
// is built from) with g++ and runs the 32 logical lanes of one warp in a loop per phase,
// with the warp shuffles, the warp scans, the batch-wide prefix and the tile stores
// replaced by their serial meaning.  Lets the bit-level logic be checked against the
// oracle on a machine without a GPU.  Never linked into libmimas_meshgen_cuda.so.
#include <stdint.h>
#include <string.h>

#include <algorithm>

#include "../../mimas_b200/csrc/mesh_core.cuh"

using namespace mm;

// the three-part store of mesh_kernel.cuh store_tile(): head / 16-byte aligned body / tail
static void store_tile(unsigned char *g, const unsigned char *sp, uint32_t bytes) {
	uint32_t head = (uint32_t)((16u - (uint32_t)((uintptr_t)g & 15u)) & 15u);
	if (head > bytes) head = bytes;
	if (head) memcpy(g, sp, 8);
	const uint32_t body = (bytes - head) & ~15u;
	if (body) memcpy(g + head, sp + head, body);
	if (bytes - head - body) memcpy(g + head + body, sp + head + body, 8);
}

static Col load_col(const uint8_t *chunk, int x, int y) {
	Col c;
	memcpy(c.w, chunk + (x * 16 + y) * 16, 16);
	return c;
}

extern "C" void emul_transpose16(uint16_t rows[16]) {
	uint32_t w[8];
	for (int k = 0; k < 8; k++) w[k] = rows[2 * k] | ((uint32_t)rows[2 * k + 1] << 16);
	transpose16(w);
	for (int k = 0; k < 8; k++) {
		rows[2 * k] = (uint16_t)w[k];
		rows[2 * k + 1] = (uint16_t)(w[k] >> 16);
	}
}

extern "C" uint32_t emul_eq_mask16(const uint8_t a[16], const uint8_t b[16]) {
	Col ca, cb;
	memcpy(ca.w, a, 16);
	memcpy(cb.w, b, 16);
	return eq_mask16(ca, cb);
}

extern "C" int emul_mesh_batch(const uint8_t *chunks, const int64_t *offs, uint32_t n, const TabEntry *tab,
                               unsigned char *out_i, uint64_t cap_i, unsigned char *out_t, uint64_t cap_t,
                               Range *ranges_i, Range *ranges_t, uint64_t *totals) {
	WarpSmem *sp = new WarpSmem();
	WarpSmem &s = *sp;
	memset(&s, 0xA5, sizeof(s));  // nothing may depend on stale shared memory
	uint8_t flags[256];
	for (int i = 0; i < 256; i++) {
		const uint32_t kind = tab[i].y >> 16;
		flags[i] = kind == 1 ? kFlagBlocky : kind == 2 ? kFlagCrossed : 0;
	}
	for (int lane = 0; lane < kLanes; lane++) init_warp_smem(s, lane);
	uint64_t run_q = 0, run_t = 0;
	for (uint32_t c = 0; c < n; c++) {
		const uint8_t *chunk = chunks + (size_t)c * 4096;
		long long o3[3] = {offs[3 * c], offs[3 * c + 1], offs[3 * c + 2]};
		for (int lane = 0; lane < kLanes; lane++) phase_pos(s, lane, o3);
		for (int lane = 0; lane < kLanes; lane++) zero_border(s, lane);
		memcpy(smem_raw(s), chunk, 4096);  // the bulk async copy of the kernel
		const uint8_t *raw = smem_raw(s);
		// A
		uint32_t anyb = 0, npl = 0;
		uint32_t crossed_cols[kLanes] = {0};
		for (int j = 0; j < 8; j++)
			for (int lane = 0; lane < kLanes; lane++) {
				const int h = lane >> 4, y = lane & 15, x = 2 * j + h;
				const Col own = load_col(chunk, x, y);
				// what the shuffles deliver: lane-1 (wraps to garbage at y = 0) and the (x-1, y) column
				Col ny = y > 0 ? load_col(chunk, x, y - 1) : own;
				Col nx;
				if (x > 0) nx = load_col(chunk, x - 1, y);
				else memset(nx.w, 0, sizeof(nx.w));
				uint32_t sum;
				const uint32_t bz = col_flags(flags, own, sum);
				smem_Bzp(s)[(x + 1) * kBzStride + y + 1] = (uint16_t)bz;
				smem_Ey(s)[32 * j + lane] = (uint16_t)eq_mask16(own, ny);
				smem_Ex(s)[32 * j + lane] = (uint16_t)eq_mask16(own, nx);
				anyb |= bz;
				crossed_cols[lane] |= ((sum >> 5) ? 1u : 0u) << j;
				npl += sum >> 5;
			}
		uint32_t anyc = 0;
		for (int lane = 0; lane < kLanes; lane++) anyc |= crossed_cols[lane];
		uint32_t nq = 0;
		if (!(anyb | anyc)) npl = 0;
		if (anyb | anyc) {
			// B
			for (int j = 0; j < 8; j++)
				for (int lane = 0; lane < kLanes; lane++) col_starts(s, tab, raw, 2 * j + (lane >> 4), lane & 15);
			// C
			for (int lane = 0; lane < kLanes; lane++) transpose_blocky(s, lane);
			uint32_t cnt[kGroups];
			for (int k = 0; k < 3; k++)
				for (int lane = 0; lane < kLanes; lane++) cnt[lane + 32 * k] = transpose_group(s, lane + 32 * k);
			for (int m = 0; m < kGroups; m++) {
				s.gpre[m] = (uint16_t)nq;
				nq += cnt[m];
			}
		}
		const uint32_t nts = npl * kPlantQuads;
		// D
		const uint64_t base_q = run_q, base_t = run_t;
		run_q += nq;
		run_t += nts;
		ranges_i[c].first_vertex = base_q * 6; ranges_i[c].vertex_count = nq * 6; ranges_i[c].reserved = 0;
		ranges_t[c].first_vertex = base_t * 6; ranges_t[c].vertex_count = nts * 6; ranges_t[c].reserved = 0;
		// E1
		if (nq && base_q + nq <= cap_i) {
			for (uint32_t t0 = 0; t0 < nq; t0 += kTileQuads) {
				const uint32_t tn = std::min<uint32_t>(kTileQuads, nq - t0);
				int m[kLanes], c2[kLanes], st[kLanes];
				for (uint32_t lane = 0; lane < tn; lane++) find_quad(s, t0 + lane, m[lane], c2[lane], st[lane]);
				unsigned char *g = out_i + (base_q + t0) * (uint64_t)kQuadBytes;
				unsigned char *stg = s.stage + ((uintptr_t)g & 8u);
				for (uint32_t lane = 0; lane < tn; lane++)
					expand_face_quad(s, tab, chunk, m[lane], c2[lane], st[lane], stg + lane * kQuadBytes);
				store_tile(g, stg, tn * kQuadBytes);
			}
		}
		// E2
		if (npl && base_t + nts <= cap_t) {
			uint16_t *Cz = smem_Cz(s), *cpre = smem_cpre(s);
			for (int j = 0; j < 8; j++)
				for (int lane = 0; lane < kLanes; lane++) {
					const int h = lane >> 4, y = lane & 15, x = 2 * j + h;
					Cz[32 * j + lane] =
					    (uint16_t)(((crossed_cols[lane] >> j) & 1u) ? col_crossed(flags, load_col(chunk, x, y)) : 0u);
				}
			uint32_t ex = 0;
			for (int col = 0; col < 256; col++) {
				cpre[col] = (uint16_t)ex;
				ex += (uint32_t)popc32(Cz[col]);
			}
			for (uint32_t t0 = 0; t0 < nts; t0 += kTileQuads) {
				const uint32_t tn = std::min<uint32_t>(kTileQuads, nts - t0);
				int cell[kLanes];
				for (uint32_t lane = 0; lane < tn; lane++) cell[lane] = find_plant(s, (t0 + lane) >> 2);
				unsigned char *g = out_t + (base_t + t0) * (uint64_t)kQuadBytes;
				unsigned char *stg = s.stage + ((uintptr_t)g & 8u);
				for (uint32_t lane = 0; lane < tn; lane++)
					expand_plant_quad(s, tab, chunk, cell[lane], (int)((t0 + lane) & 3u), stg + lane * kQuadBytes);
				store_tile(g, stg, tn * kQuadBytes);
			}
		}
	}
	totals[0] = run_q;
	totals[1] = run_t;
	delete sp;
	return 0;
}
