// TEST PROGRAM for include/mimas_meshgen.hpp, the C++ mirror of the mimas-meshgen crate API.
//   host_mirror_test selfcheck        host logic only (id allocation order, table flattening, loud failure
//                                     without a CUDA device)
//   host_mirror_test cases <file>     mesh_for_chunk / mesh_batch / push_block on a B200 against expected bytes
//                                     written by tests/test_cxx_host.py (known-answer vectors + oracle output)
//   host_mirror_test latency <file>   the reference's call pattern from a compiled host: one mesh_for_chunk call per
//                                     chunk of the file (client.rs:259-267), microseconds per call incl. the Vec copies
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>

#include "../../include/mimas_meshgen.hpp"

using namespace mimas_meshgen;

#define CHECK(c)                                                                  \
	do {                                                                          \
		if (!(c)) {                                                               \
			std::fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #c); \
			return 1;                                                             \
		}                                                                         \
	} while (0)

static int selfcheck() {
	// lib.rs:87-104 + assets.rs:145-201: fallback Colored -> ids 0 (top/bottom), 1 (sides); then blocks in order
	GameParamsHdl hdl;
	hdl.block_params.resize(6);
	hdl.block_params[1].draw_style = DrawStyle::texture("water");                       // id 2
	hdl.block_params[2].draw_style = DrawStyle::sides_top("tree", "tree_top");           // s 3, t/b 4
	hdl.block_params[3].draw_style = DrawStyle::sides_top_bottom("gs", "gt", "gb");      // 5, 6, 7
	hdl.block_params[4].draw_style = DrawStyle::crossed("sapling");                      // 8
	hdl.block_params[5].draw_style = DrawStyle::colored({1, 0, 0, 1});                   // t/b 9, sides 10
	hdl.block_params[1].inv_texture = "inv_water";                                       // 11
	HeadlessAssets ob;
	const TextureIdCache cache = TextureIdCache::from_hdl(hdl, ob);
	CHECK(ob.count == 12);
	CHECK(!cache.get_bl_tex_ids(0).has_value());
	CHECK(cache.get_bl_tex_ids(1)->id_sides.v == 2 && cache.get_bl_tex_ids(1)->id_top.v == 2);
	CHECK(cache.get_bl_tex_ids(2)->id_sides.v == 3 && cache.get_bl_tex_ids(2)->id_top.v == 4 && cache.get_bl_tex_ids(2)->id_bottom.v == 4);
	CHECK(cache.get_bl_tex_ids(3)->id_sides.v == 5 && cache.get_bl_tex_ids(3)->id_top.v == 6 && cache.get_bl_tex_ids(3)->id_bottom.v == 7);
	CHECK(!cache.get_bl_tex_ids(4).has_value());                       // crossed: not blocky (lib.rs:92-94)
	CHECK(cache.get_mesh_draw_style(4)->kind == MeshDrawStyle::Crossed && cache.get_mesh_draw_style(4)->crossed_id.v == 8);
	CHECK(cache.get_bl_tex_ids(5)->id_top.v == 9 && cache.get_bl_tex_ids(5)->id_sides.v == 10);
	CHECK(cache.get_bl_tex_ids(200)->id_sides.v == 1 && cache.get_bl_tex_ids(200)->id_top.v == 0);  // fallback, lib.rs:106-108
	CHECK(!cache.get_mesh_draw_style(200).has_value());                // lib.rs:110-113
	CHECK(cache.get_inv_texture_id(1)->v == 11 && !cache.get_inv_texture_id(2).has_value());
	const auto t = cache.to_table();
	CHECK(t[0].kind == MM_KIND_NONE && t[4].kind == MM_KIND_CROSSED && t[4].id_crossed == 8);
	CHECK(t[3].kind == MM_KIND_BLOCKY && t[3].id_bottom == 7);
	CHECK(t[255].kind == MM_KIND_BLOCKY && t[255].id_sides == 1 && t[255].id_top == 0 && t[255].id_bottom == 0);
	static_assert(sizeof(Vertex) == 36, "Vertex layout");
	try {
		MapChunkData chunk;
		chunk.set_blk(0, 0, 0, 1);
		const ChunkMesh m = mesh_for_chunk({0, 0, 0}, chunk, cache);
		CHECK(m.intransparent.size() == 36 && m.transparent.empty());  // one cube: six quads
		std::printf("selfcheck ok (CUDA device present: one cube -> %zu vertices)\n", m.intransparent.size());
	} catch (const std::runtime_error &e) {
		std::printf("selfcheck ok (no device: %s)\n", e.what());
	}
	return 0;
}

template <class T> static T rd(std::ifstream &f) {
	T v;
	f.read(reinterpret_cast<char *>(&v), sizeof(v));
	return v;
}

static int cases(const char *path) {
	std::ifstream f(path, std::ios::binary);
	CHECK(f.good());
	CHECK(rd<uint32_t>(f) == 0x54434D4Du);
	const uint32_t n_rows = rd<uint32_t>(f);
	uint16_t fb[4];
	f.read(reinterpret_cast<char *>(fb), 8);
	std::vector<std::optional<MeshDrawStyle>> styles;
	for (uint32_t i = 0; i < n_rows; i++) {
		uint16_t r[5];
		f.read(reinterpret_cast<char *>(r), 10);
		if (r[0] == 1) styles.push_back(MeshDrawStyle::make_blocky(BlockTextureIds::make({r[1]}, {r[2]}, {r[3]})));
		else if (r[0] == 2) styles.push_back(MeshDrawStyle::make_crossed({r[4]}));
		else styles.push_back(std::nullopt);
	}
	const TextureIdCache cache(BlockTextureIds::make({fb[0]}, {fb[1]}, {fb[2]}), styles);
	const uint32_t n = rd<uint32_t>(f);
	std::vector<std::array<isize, 3>> offs(n);
	std::vector<MapChunkData> chunks(n);
	std::vector<std::vector<char>> want_i(n), want_t(n);
	for (uint32_t i = 0; i < n; i++) {
		f.read(reinterpret_cast<char *>(offs[i].data()), 24);
		f.read(reinterpret_cast<char *>(chunks[i].ids.data()), 4096);
		const uint64_t ni = rd<uint64_t>(f), nt = rd<uint64_t>(f);
		want_i[i].resize(ni * 36);
		want_t[i].resize(nt * 36);
		f.read(want_i[i].data(), (std::streamsize)want_i[i].size());
		f.read(want_t[i].data(), (std::streamsize)want_t[i].size());
	}
	CHECK(f.good());
	auto same = [](const std::vector<Vertex> &got, const std::vector<char> &want) {
		return got.size() * 36 == want.size() && (want.empty() || std::memcmp(got.data(), want.data(), want.size()) == 0);
	};
	for (uint32_t i = 0; i < n; i++) {  // the crate's entry point, one chunk per call
		const ChunkMesh m = mesh_for_chunk(offs[i], chunks[i], cache);
		if (!same(m.intransparent, want_i[i]) || !same(m.transparent, want_t[i])) {
			std::fprintf(stderr, "mesh_for_chunk: case %u differs\n", i);
			return 1;
		}
	}
	Mesher mesher(0);  // batch submission
	const std::vector<ChunkMesh> all = mesher.mesh_batch(offs, chunks, cache);
	for (uint32_t i = 0; i < n; i++)
		if (!same(all[i].intransparent, want_i[i]) || !same(all[i].transparent, want_t[i])) {
			std::fprintf(stderr, "mesh_batch: case %u differs\n", i);
			return 1;
		}
	// push_block with nothing blocked == the six faces of a lone cube at the origin with siz 1 (lib.rs:183-219)
	std::vector<Vertex> r;
	push_block(r, {0.f, 0.f, 0.f}, BlockTextureIds::make({6}, {6}, {6}), 1.0f, [](std::array<isize, 3>) { return false; });
	CHECK(r.size() == 36);
	std::printf("cases ok: %u chunks bit-exact through mesh_for_chunk and mesh_batch\n", n);
	return 0;
}

static int latency(const char *path) {
	std::ifstream f(path, std::ios::binary);
	CHECK(f.good());
	CHECK(rd<uint32_t>(f) == 0x54434D4Du);
	const uint32_t n_rows = rd<uint32_t>(f);
	uint16_t fb[4];
	f.read(reinterpret_cast<char *>(fb), 8);
	std::vector<std::optional<MeshDrawStyle>> styles;
	for (uint32_t i = 0; i < n_rows; i++) {
		uint16_t r[5];
		f.read(reinterpret_cast<char *>(r), 10);
		if (r[0] == 1) styles.push_back(MeshDrawStyle::make_blocky(BlockTextureIds::make({r[1]}, {r[2]}, {r[3]})));
		else if (r[0] == 2) styles.push_back(MeshDrawStyle::make_crossed({r[4]}));
		else styles.push_back(std::nullopt);
	}
	const TextureIdCache cache(BlockTextureIds::make({fb[0]}, {fb[1]}, {fb[2]}), styles);
	const uint32_t n = rd<uint32_t>(f);
	std::vector<std::array<isize, 3>> offs(n);
	std::vector<MapChunkData> chunks(n);
	for (uint32_t i = 0; i < n; i++) {
		f.read(reinterpret_cast<char *>(offs[i].data()), 24);
		f.read(reinterpret_cast<char *>(chunks[i].ids.data()), 4096);
		const uint64_t ni = rd<uint64_t>(f), nt = rd<uint64_t>(f);
		f.seekg((std::streamoff)((ni + nt) * 36), std::ios::cur);
	}
	CHECK(f.good());
	for (uint32_t i = 0; i < n; i++) {
		for (int w = 0; w < 20; w++) (void)mesh_for_chunk(offs[i], chunks[i], cache);
		std::vector<double> us(300);
		size_t verts = 0;
		for (double &u : us) {
			const auto t0 = std::chrono::steady_clock::now();
			const ChunkMesh m = mesh_for_chunk(offs[i], chunks[i], cache);
			u = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
			verts = m.intransparent.size() + m.transparent.size();
		}
		std::sort(us.begin(), us.end());
		std::printf("{\"chunk\": %u, \"offs\": [%lld, %lld, %lld], \"vertices\": %zu, \"us_median\": %.2f, \"us_p10\": %.2f, \"us_p90\": %.2f}\n", i,
		            (long long)offs[i][0], (long long)offs[i][1], (long long)offs[i][2], verts, us[150], us[30], us[270]);
	}
	return 0;
}

int main(int argc, char **argv) {
	try {
		if (argc >= 2 && std::string(argv[1]) == "selfcheck") return selfcheck();
		if (argc >= 3 && std::string(argv[1]) == "cases") return cases(argv[2]);
		if (argc >= 3 && std::string(argv[1]) == "latency") return latency(argv[2]);
	} catch (const std::exception &e) {
		std::fprintf(stderr, "exception: %s\n", e.what());
		return 2;
	}
	std::fprintf(stderr, "usage: host_mirror_test selfcheck | cases <file>\n");
	return 64;
}
