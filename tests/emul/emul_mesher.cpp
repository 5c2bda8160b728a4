// TEST INFRASTRUCTURE -- host emulation of the mesher kernel's control flow.
//
// Compiles mimas_b200/csrc/mesh_core.cuh (the per-lane phase functions the CUDA kernel
// is built from) with g++ and runs the threads of one analysing CTA / the lanes of one expanding warp in a loop per phase,
// with the warp shuffles, the warp scans, the batch-wide prefix and the tile stores
// replaced by their serial meaning.  Lets the bit-level logic be checked against the
// oracle on a machine without a GPU.  Never linked into libmimas_meshgen_cuda.so.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

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

template <bool HALO>
static int mesh_batch_t(const uint8_t *chunks, const int64_t *offs, uint32_t n, const TabEntry *tab,
                        unsigned char *out_i, uint64_t cap_i, unsigned char *out_t, uint64_t cap_t,
                        Range *ranges_i, Range *ranges_t, uint64_t *totals, const uint8_t *halo) {
	ChunkSmem *sp = new ChunkSmem();
	ChunkSmem &s = *sp;
	memset(&s, 0xA5, sizeof(s));  // nothing may depend on stale shared memory
	uint8_t flags[256];
	build_flags(tab, flags);
	for (int lane = 0; lane < kLanes; lane++) init_warp_smem(s, lane);
	// ---- analyse kernel: one CTA per chunk -> counts + descriptors ----------------------
	std::vector<uint32_t> desc_i((size_t)n * kDescStrideI, 0xDEADBEEFu), desc_t((size_t)n * kDescStrideT, 0xDEADBEEFu);
	std::vector<uint32_t> cnt_i(n), cnt_t(n);
	for (uint32_t c = 0; c < n; c++) {
		const uint8_t *chunk = chunks + (size_t)c * 4096;
		memcpy(smem_raw(s), chunk, 4096);  // the bulk async copy of the kernel
		const uint8_t *raw = smem_raw(s);
		// A
		uint32_t amb = 0;
		uint32_t crossed_cols[kLanes * 4] = {0};
		bool uniform = true;
		for (int i = 1; i < 4096; i++) uniform = uniform && raw[i] == raw[0];
		if (uniform && (flags[raw[0]] & (kFlagBlocky | kFlagCrossed)) == 0) {  // the kernel's all-air exit
			cnt_i[c] = 0;
			cnt_t[c] = 0;
			continue;
		}
		if (HALO)
			for (int tid = 0; tid < 96; tid++) halo_row(s, flags, load_col(halo + (size_t)c * 1536 + tid * 16), tid);
		for (int j = 0; j < 2; j++)
			for (int tid = 0; tid < 128; tid++) {
				const int col = 128 * j + tid;
				uint32_t orf;
				const uint32_t bz = col_flags(flags, load_col(raw + col * 16), orf);
				smem_Bzp(s)[((col >> 4) + 1) * kBzStride + (col & 15) + 1] = (uint16_t)bz;
				crossed_cols[tid] |= ((orf >> 1) & 1u) << j;
				amb |= orf;
			}
		uint32_t nq = 0, npl = 0;
		uint32_t *di = desc_i.data() + (size_t)c * kDescStrideI;
		if (amb & kFlagBlocky) {
			// B
			for (int col = 0; col < 256; col++) col_starts<HALO>(s, tab, raw, col >> 4, col & 15, amb);
			// C
			for (int lane = 0; lane < kLanes; lane++) transpose_blocky(s, lane);
			if (HALO)
				for (int job = 0; job < 4; job++) transpose_halo(s, job);
			uint32_t run = 0;
			for (int m = 0; m < kGroups; m++) s.gcnt[m] = transpose_group(s, m);
			for (int m = 0; m < kGroups; m++) {  // exclusive scan, quads (low half) and words (high half) at once
				const uint32_t v = s.gcnt[m];
				s.gcnt[m] = run;
				run += v;
			}
			nq = run & 0xffffu;
			const uint32_t nwords = run >> 16;
			for (int m = 0; m < kGroups; m++) list_group_words(s, m, s.gcnt[m]);
			// D
			for (uint32_t i = 0; i < nwords; i++) {
				const uint32_t e = smem_wl(s)[i];
				const uint32_t widx = e & 1023u;
				emit_word_descs<HALO>(s, widx, reinterpret_cast<const uint32_t *>(s.S)[widx], di + (e >> 10));
			}
		}
		if (amb & kFlagCrossed) {
			uint16_t *Cz = smem_Cz(s);
			for (int j = 0; j < 2; j++)
				for (int tid = 0; tid < 128; tid++) {
					const int col = 128 * j + tid;
					Cz[col] = (uint16_t)(((crossed_cols[tid] >> j) & 1u) ? col_crossed(flags, load_col(raw + col * 16)) : 0u);
				}
			for (int lane = 0; lane < kLanes; lane++) {
				const uint32_t mine = count_plants8(s, lane);
				if (mine) emit_plant_descs(s, lane, desc_t.data() + (size_t)c * kDescStrideT + npl);
				npl += mine;
			}
		}
		cnt_i[c] = nq;
		cnt_t[c] = npl * kPlantQuads;
	}
	delete sp;
	// ---- expand kernel: scan, then tiles of 32 quads of the packed streams -----------------
	std::vector<uint32_t> pre_i(n + 1), pre_t(n + 1);
	pre_i[0] = pre_t[0] = 0;
	for (uint32_t c = 0; c < n; c++) {
		pre_i[c + 1] = pre_i[c] + cnt_i[c];
		pre_t[c + 1] = pre_t[c] + cnt_t[c];
		ranges_i[c].first_vertex = (uint64_t)pre_i[c] * 6; ranges_i[c].vertex_count = cnt_i[c] * 6; ranges_i[c].reserved = 0;
		ranges_t[c].first_vertex = (uint64_t)pre_t[c] * 6; ranges_t[c].vertex_count = cnt_t[c] * 6; ranges_t[c].reserved = 0;
	}
	totals[0] = pre_i[n];
	totals[1] = pre_t[n];
	alignas(16) unsigned char stage[kStageBytes + 16];
	for (int face = 1; face >= 0; face--) {
		const std::vector<uint32_t> &pre = face ? pre_i : pre_t;
		const uint32_t tot = pre[n];
		unsigned char *out = face ? out_i : out_t;
		if (tot > (face ? cap_i : cap_t)) continue;
		for (uint32_t q0 = 0; q0 < tot; q0 += kTileQuads) {
			const uint32_t tn = std::min<uint32_t>(kTileQuads, tot - q0);
			unsigned char *g = out + (uint64_t)q0 * kQuadBytes;
			unsigned char *stg = stage + ((uintptr_t)g & 8u);
			for (uint32_t lane = 0; lane < tn; lane++) {
				const uint32_t q = q0 + lane;
				const uint32_t c = (uint32_t)(std::upper_bound(pre.begin(), pre.begin() + n, q) - pre.begin()) - 1;  // largest c with pre[c] <= q
				const uint32_t r = q - pre[c];
				const long long ox = offs[3 * c], oy = offs[3 * c + 1], oz = offs[3 * c + 2];
				if (face) expand_face_desc(tab, ox, oy, oz, desc_i[(size_t)c * kDescStrideI + r], stg + lane * kQuadBytes);
				else expand_plant_desc(tab, ox, oy, oz, desc_t[(size_t)c * kDescStrideT + (r >> 2)], (int)(q & 3u), stg + lane * kQuadBytes);
			}
			store_tile(g, stg, tn * kQuadBytes);
		}
	}
	return 0;
}

extern "C" int emul_mesh_batch(const uint8_t *chunks, const int64_t *offs, uint32_t n, const TabEntry *tab,
                               unsigned char *out_i, uint64_t cap_i, unsigned char *out_t, uint64_t cap_t,
                               Range *ranges_i, Range *ranges_t, uint64_t *totals) {
	return mesh_batch_t<false>(chunks, offs, n, tab, out_i, cap_i, out_t, cap_t, ranges_i, ranges_t, totals, nullptr);
}

// halo: n * 1536 block ids (see halo_row in mesh_core.cuh), 16-byte aligned
extern "C" int emul_mesh_batch_halo(const uint8_t *chunks, const int64_t *offs, uint32_t n, const TabEntry *tab,
                                    unsigned char *out_i, uint64_t cap_i, unsigned char *out_t, uint64_t cap_t,
                                    Range *ranges_i, Range *ranges_t, uint64_t *totals, const uint8_t *halo) {
	return mesh_batch_t<true>(chunks, offs, n, tab, out_i, cap_i, out_t, cap_t, ranges_i, ranges_t, totals, halo);
}
