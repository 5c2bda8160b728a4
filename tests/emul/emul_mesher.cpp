// TEST INFRASTRUCTURE -- host emulation of the mesher kernel's control flow.
//
// Compiles mimas_b200/csrc/mesh_core.cuh (the per-thread phase functions the CUDA kernel
// is built from) with g++ and runs the 128 logical threads of one CTA in a loop, with the
// CTA-wide scan, the look-back prefix and the tile stores replaced by their serial
// meaning.  Lets the bit-level logic be checked against the oracle on a machine without
// a GPU.  Never linked into libmimas_meshgen_cuda.so.
#include <stdint.h>
#include <string.h>

#include <algorithm>

#include "../../mimas_b200/csrc/mesh_core.cuh"

using namespace mm;

static uint32_t serial_exclusive_scan(uint16_t *a, int n) {
	uint32_t run = 0;
	for (int i = 0; i < n; i++) {
		const uint32_t v = a[i];
		a[i] = (uint16_t)run;
		run += v;
	}
	return run;
}

// the three-part store of mesh_kernel.cuh store_tile(): head / 16-byte aligned body / tail
static void store_tile(unsigned char *g, const unsigned char *sp, uint32_t bytes) {
	uint32_t head = (uint32_t)((16u - (uint32_t)((uintptr_t)g & 15u)) & 15u);
	if (head > bytes) head = bytes;
	if (head) memcpy(g, sp, 8);
	const uint32_t body = (bytes - head) & ~15u;
	if (body) memcpy(g + head, sp + head, body);
	if (bytes - head - body) memcpy(g + head + body, sp + head + body, 8);
}

extern "C" int emul_mesh_batch(const uint8_t *chunks, const int64_t *offs, uint32_t n, const TabEntry *tab,
                               unsigned char *out_i, uint64_t cap_i, unsigned char *out_t, uint64_t cap_t,
                               Range *ranges_i, Range *ranges_t, uint64_t *totals) {
	Smem *sp = new Smem();
	Smem &s = *sp;
	memcpy(s.tab, tab, sizeof(s.tab));
	uint64_t run_q = 0, run_t = 0;
	for (uint32_t c = 0; c < n; c++) {
		memcpy(s.raw, chunks + (size_t)c * 4096, 4096);
		long long o3[3] = {offs[3 * c], offs[3 * c + 1], offs[3 * c + 2]};
		for (int tid = 0; tid < kThreads; tid++) phase_pos(s, tid, o3);
		bool uniform = true;
		for (int i = 1; i < 4096; i++) uniform &= s.raw[i] == s.raw[0];
		uint32_t nq = 0, npl = 0;
		const bool trivial = uniform && (s.tab[s.raw[0]].y >> 16) == 0u;
		if (!trivial) {
			RowRegs g[kRows];
			uint32_t cr = 0;
			for (int r = 0; r < kRows; r++) g[r] = phase_rows(s, r, cr);
			for (int r = 0; r < kRows; r++) phase_vis(s, r, g[r]);
			if (cr)
				for (int col = 0; col < kRows; col++) phase_crossed_cols(s, col);
			nq = serial_exclusive_scan(s.pre, kPasses * kRows);
			if (cr) npl = serial_exclusive_scan(s.cpre, kRows);
		}
		const uint32_t nts = npl * kPlantQuads;
		const uint64_t base_q = run_q, base_t = run_t;
		run_q += nq;
		run_t += nts;
		ranges_i[c].first_vertex = base_q * 6; ranges_i[c].vertex_count = nq * 6; ranges_i[c].reserved = 0;
		ranges_t[c].first_vertex = base_t * 6; ranges_t[c].vertex_count = nts * 6; ranges_t[c].reserved = 0;
		if (nq && base_q + nq <= cap_i) {
			for (uint32_t w0 = 0; w0 < nq; w0 += kDescCap) {
				for (int r = 0; r < kRows; r++) phase_desc_row(s, r, w0);
				const uint32_t wend = std::min<uint32_t>(nq, w0 + kDescCap);
				for (uint32_t t0 = w0; t0 < wend; t0 += kTileQuads) {
					const uint32_t tn = std::min<uint32_t>(kTileQuads, wend - t0);
					unsigned char *g = out_i + (base_q + t0) * (uint64_t)kQuadBytes;
					unsigned char *st = s.stage + ((uintptr_t)g & 8u);
					for (uint32_t tid = 0; tid < tn; tid++)
						expand_face_quad(s, s.desc[t0 - w0 + tid], st + tid * kQuadBytes);
					store_tile(g, st, tn * kQuadBytes);
				}
			}
		}
		if (npl && base_t + nts <= cap_t) {
			for (uint32_t w0 = 0; w0 < npl; w0 += kDescCap) {
				for (int col = 0; col < kRows; col++) phase_desc_col(s, col, w0);
				const uint32_t qend = std::min<uint32_t>(npl, w0 + kDescCap) * kPlantQuads;
				for (uint32_t t0 = w0 * kPlantQuads; t0 < qend; t0 += kTileQuads) {
					const uint32_t tn = std::min<uint32_t>(kTileQuads, qend - t0);
					unsigned char *g = out_t + (base_t + t0) * (uint64_t)kQuadBytes;
					unsigned char *st = s.stage + ((uintptr_t)g & 8u);
					for (uint32_t tid = 0; tid < tn; tid++) {
						const uint32_t q = t0 + tid;
						expand_plant_quad(s, s.desc[(q >> 2) - w0], (int)(q & 3u), st + tid * kQuadBytes);
					}
					store_tile(g, st, tn * kQuadBytes);
				}
			}
		}
	}
	totals[0] = run_q;
	totals[1] = run_t;
	delete sp;
	return 0;
}
