// mimas_meshgen.hpp -- C++ host side above the C ABI (include/mimas_meshgen_cuda.h), mirroring the
// public API of the reference crate mimas-meshgen (/root/reference/mimas-meshgen/lib.rs) name for
// name: Vertex (:18-24), TextureId / BlockTextureIds / MeshDrawStyle (:28-71), TextureIdCache and
// TextureIdObtainer (:73-118), push_block (:183-219), ChunkMesh (:226-229), mesh_for_chunk (:231-232).
// The reference is Rust; this image has no Rust toolchain, so the drop-in host code is written in
// C++ (the Rust shim a maintainer would add is in INTEGRATION.md).  Header only; link with
// -lmimas_meshgen_cuda.  There is no CPU mesher behind it: without a CUDA device every call throws.
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <functional>
#include <optional>
#include <stdexcept>
#include <string>
#include <variant>
#include <vector>

#include "mimas_meshgen_cuda.h"

namespace mimas_meshgen {

using isize = int64_t;                     // Rust isize on the 64-bit targets mimas builds for
constexpr size_t CHUNKSIZE = MM_CHUNKSIZE;  // mimas-common/map.rs:11

// lib.rs:18-24; #[repr(C)] in declared order
using Vertex = mm_vertex;

// mimas-common/map.rs:14,56-61: MapBlock(u8), MapChunkData([MapBlock; 4096], metadata) -- the mesher
// only reads the ids, index x*256 + y*16 + z (map.rs:83-86)
struct MapChunkData {
	std::array<uint8_t, MM_CHUNK_BYTES> ids{};
	uint8_t get_blk(size_t x, size_t y, size_t z) const { return ids[x * 256 + y * 16 + z]; }
	void set_blk(size_t x, size_t y, size_t z, uint8_t id) { ids[x * 256 + y * 16 + z] = id; }
};

// lib.rs:28-29
struct TextureId {
	uint16_t v = 0;
	bool operator==(const TextureId &o) const { return v == o.v; }
};

// lib.rs:31-58
struct BlockTextureIds {
	TextureId id_sides, id_top, id_bottom;
	static BlockTextureIds make(TextureId s, TextureId t, TextureId b) { return {s, t, b}; }   // new()
	static BlockTextureIds new_tb(TextureId top_bottom, TextureId sides) { return {sides, top_bottom, top_bottom}; }
	static BlockTextureIds uniform(TextureId id) { return {id, id, id}; }
};

// lib.rs:60-71
struct MeshDrawStyle {
	enum Kind { Blocky, Crossed } kind = Blocky;
	BlockTextureIds blocky_ids{};
	TextureId crossed_id{};
	static MeshDrawStyle make_blocky(BlockTextureIds b) { return {Blocky, b, {}}; }
	static MeshDrawStyle make_crossed(TextureId t) { return {Crossed, {}, t}; }
	std::optional<BlockTextureIds> blocky() const {
		return kind == Blocky ? std::optional<BlockTextureIds>(blocky_ids) : std::nullopt;
	}
};

// mimas-common/game_params.rs:23-29
struct DrawStyle {
	enum Kind { Colored, Crossed, Texture, TextureSidesTop, TextureSidesTopBottom } kind = Colored;
	std::array<float, 4> color{0, 0, 0, 1};
	std::array<std::string, 3> paths;
	static DrawStyle colored(std::array<float, 4> c) { DrawStyle d; d.kind = Colored; d.color = c; return d; }
	static DrawStyle crossed(std::string p) { DrawStyle d; d.kind = Crossed; d.paths[0] = std::move(p); return d; }
	static DrawStyle texture(std::string p) { DrawStyle d; d.kind = Texture; d.paths[0] = std::move(p); return d; }
	static DrawStyle sides_top(std::string s, std::string tb) {
		DrawStyle d; d.kind = TextureSidesTop; d.paths[0] = std::move(s); d.paths[1] = std::move(tb); return d;
	}
	static DrawStyle sides_top_bottom(std::string s, std::string t, std::string b) {
		DrawStyle d; d.kind = TextureSidesTopBottom; d.paths = {std::move(s), std::move(t), std::move(b)}; return d;
	}
};

// the two fields of BlockParams the mesher's table is built from (game_params.rs:178-195)
struct BlockParams {
	std::optional<DrawStyle> draw_style;
	std::optional<std::string> inv_texture;
};
struct GameParamsHdl {
	std::vector<BlockParams> block_params;
};

// lib.rs:81-84
struct TextureIdObtainer {
	virtual ~TextureIdObtainer() = default;
	virtual MeshDrawStyle mesh_draw_st(const DrawStyle &ds) = 0;
	virtual TextureId inv_texture_id(const std::string &path) = 0;
};

// Texture-id allocation of mimas-client/assets.rs:140-201 without loading any pixels: every
// add_asset hands out the next index (assets.rs:140-144).
struct HeadlessAssets : TextureIdObtainer {
	uint32_t count = 0;
	TextureId add_asset() {
		if (count > 0xFFFF) throw std::overflow_error("TextureId is a u16");
		return TextureId{(uint16_t)count++};
	}
	MeshDrawStyle mesh_draw_st(const DrawStyle &ds) override {
		switch (ds.kind) {
		case DrawStyle::Colored: {  // assets.rs:148-152
			const TextureId id = add_asset(), id_h = add_asset();
			return MeshDrawStyle::make_blocky(BlockTextureIds::new_tb(id, id_h));
		}
		case DrawStyle::Crossed: return MeshDrawStyle::make_crossed(add_asset());                        // :153-158
		case DrawStyle::Texture: return MeshDrawStyle::make_blocky(BlockTextureIds::uniform(add_asset())); // :159-164
		case DrawStyle::TextureSidesTop: {  // :165-173
			const TextureId s = add_asset(), tb = add_asset();
			return MeshDrawStyle::make_blocky(BlockTextureIds::new_tb(tb, s));
		}
		default: {  // :174-186
			const TextureId s = add_asset(), t = add_asset(), b = add_asset();
			return MeshDrawStyle::make_blocky(BlockTextureIds::make(s, t, b));
		}
		}
	}
	TextureId inv_texture_id(const std::string &) override { return add_asset(); }  // assets.rs:189-194
};

// lib.rs:73-118
class TextureIdCache {
public:
	TextureIdCache(BlockTextureIds fallback, std::vector<std::optional<MeshDrawStyle>> styles,
	               std::vector<std::optional<TextureId>> inv = {})
	    : fallback_id(fallback), mesh_draw_styles(std::move(styles)), inv_texture_ids(std::move(inv)) {
		if (mesh_draw_styles.size() > MM_TABLE_ENTRIES) throw std::invalid_argument("MapBlock ids are u8");
		for (const auto &s : mesh_draw_styles) block_texture_ids.push_back(s ? s->blocky() : std::nullopt);
	}
	// lib.rs:87-104: fallback first, then every block's draw style in id order, then every inv texture
	static TextureIdCache from_hdl(const GameParamsHdl &hdl, TextureIdObtainer &ob) {
		const BlockTextureIds fb = ob.mesh_draw_st(DrawStyle::colored({0.0f, 0.0f, 0.0f, 1.0f})).blocky().value();
		std::vector<std::optional<MeshDrawStyle>> styles;
		for (const auto &p : hdl.block_params)
			styles.push_back(p.draw_style ? std::optional<MeshDrawStyle>(ob.mesh_draw_st(*p.draw_style)) : std::nullopt);
		std::vector<std::optional<TextureId>> inv;
		for (const auto &p : hdl.block_params)
			inv.push_back(p.inv_texture ? std::optional<TextureId>(ob.inv_texture_id(*p.inv_texture)) : std::nullopt);
		return TextureIdCache(fb, std::move(styles), std::move(inv));
	}
	std::optional<BlockTextureIds> get_bl_tex_ids(uint8_t bl) const {  // lib.rs:105-109
		return bl < block_texture_ids.size() ? block_texture_ids[bl] : std::optional<BlockTextureIds>(fallback_id);
	}
	std::optional<MeshDrawStyle> get_mesh_draw_style(uint8_t bl) const {  // lib.rs:110-113
		return bl < mesh_draw_styles.size() ? mesh_draw_styles[bl] : std::nullopt;
	}
	std::optional<TextureId> get_inv_texture_id(uint8_t bl) const {  // lib.rs:114-117
		return bl < inv_texture_ids.size() ? inv_texture_ids[bl] : std::nullopt;
	}
	// the 256-row device table (mm_table_from_cache): unknown ids = blocky fallback, never crossed
	std::array<mm_block_entry, MM_TABLE_ENTRIES> to_table() const {
		std::vector<mm_block_entry> known(mesh_draw_styles.size());
		for (size_t i = 0; i < known.size(); i++) {
			mm_block_entry e;
			std::memset(&e, 0, sizeof(e));
			if (const auto &s = mesh_draw_styles[i]) {
				if (s->kind == MeshDrawStyle::Blocky) {
					e.kind = MM_KIND_BLOCKY;
					e.id_sides = s->blocky_ids.id_sides.v; e.id_top = s->blocky_ids.id_top.v; e.id_bottom = s->blocky_ids.id_bottom.v;
				} else {
					e.kind = MM_KIND_CROSSED;
					e.id_crossed = s->crossed_id.v;
				}
			}
			known[i] = e;
		}
		const uint16_t fb[3] = {fallback_id.id_sides.v, fallback_id.id_top.v, fallback_id.id_bottom.v};
		std::array<mm_block_entry, MM_TABLE_ENTRIES> out;
		if (mm_table_from_cache(fb, known.data(), (uint32_t)known.size(), out.data()) != MM_OK)
			throw std::invalid_argument("TextureIdCache does not flatten into a block table");
		return out;
	}

private:
	BlockTextureIds fallback_id;
	std::vector<std::optional<BlockTextureIds>> block_texture_ids;
	std::vector<std::optional<MeshDrawStyle>> mesh_draw_styles;
	std::vector<std::optional<TextureId>> inv_texture_ids;
};

// lib.rs:226-229
struct ChunkMesh {
	std::vector<Vertex> intransparent;
	std::vector<Vertex> transparent;
};

// The state the reference's meshgen thread owns (mimas-client/client.rs:258-267): one per thread and GPU.
class Mesher {
public:
	explicit Mesher(int device = 0) {
		const int rc = mm_ctx_create(device, &ctx_);
		if (rc != MM_OK) throw std::runtime_error("mimas_meshgen: no usable CUDA device (status " + std::to_string(rc) +
		                                          "); this mesher has no CPU fallback");
	}
	~Mesher() { if (ctx_) mm_ctx_destroy(ctx_); }
	Mesher(const Mesher &) = delete;
	Mesher &operator=(const Mesher &) = delete;

	void set_cache(const TextureIdCache &cache) {
		const auto t = cache.to_table();
		if (have_table_ && std::memcmp(t.data(), table_.data(), sizeof(table_)) == 0) return;
		check(mm_set_block_table(ctx_, t.data()));
		table_ = t;
		have_table_ = true;
	}
	ChunkMesh mesh_for_chunk(const std::array<isize, 3> &offs, const MapChunkData &chunk, const TextureIdCache &cache) {
		set_cache(cache);
		const mm_vertex *pi = nullptr, *pt = nullptr;
		uint64_t ni = 0, nt = 0;
		check(mm_mesh_for_chunk(ctx_, offs.data(), chunk.ids.data(), &pi, &ni, &pt, &nt));
		return ChunkMesh{std::vector<Vertex>(pi, pi + ni), std::vector<Vertex>(pt, pt + nt)};
	}
	// batch submission (north_star: mapgen chunks of mimas-server): n independent chunks, one call
	std::vector<ChunkMesh> mesh_batch(const std::vector<std::array<isize, 3>> &offs, const std::vector<MapChunkData> &chunks,
	                                  const TextureIdCache &cache) {
		if (offs.size() != chunks.size()) throw std::invalid_argument("offs and chunks differ in length");
		set_cache(cache);
		static_assert(sizeof(MapChunkData) == MM_CHUNK_BYTES && sizeof(std::array<isize, 3>) == 24, "packed inputs");
		mm_host_mesh hm;
		check(mm_mesh_batch_host(ctx_, reinterpret_cast<const uint8_t *>(chunks.data()),
		                         reinterpret_cast<const int64_t *>(offs.data()), (uint32_t)chunks.size(), &hm));
		std::vector<ChunkMesh> out(chunks.size());
		for (size_t i = 0; i < out.size(); i++) {
			const mm_range &ri = hm.ranges_i[i], &rt = hm.ranges_t[i];
			out[i].intransparent.assign(hm.intransparent + ri.first_vertex, hm.intransparent + ri.first_vertex + ri.vertex_count);
			out[i].transparent.assign(hm.transparent + rt.first_vertex, hm.transparent + rt.first_vertex + rt.vertex_count);
		}
		return out;
	}
	// lib.rs:183-219
	void push_block(std::vector<Vertex> &r, const std::array<float, 3> &pos, BlockTextureIds ids, float siz,
	                const std::function<bool(std::array<isize, 3>)> &blocked) {
		static const isize order[6][3] = {{0, 0, -1}, {0, -1, 0}, {-1, 0, 0}, {0, 0, 1}, {0, 1, 0}, {1, 0, 0}};
		mm_cube c;
		std::memset(&c, 0, sizeof(c));
		c.pos[0] = pos[0]; c.pos[1] = pos[1]; c.pos[2] = pos[2];
		c.siz = siz;
		c.id_sides = ids.id_sides.v; c.id_top = ids.id_top.v; c.id_bottom = ids.id_bottom.v;
		for (int i = 0; i < 6; i++)
			if (blocked({order[i][0], order[i][1], order[i][2]})) c.blocked_mask |= (uint16_t)(1u << i);
		mm_vertex buf[36];
		uint64_t n = 0;
		check(mm_push_blocks(ctx_, &c, 1, buf, &n));
		r.insert(r.end(), buf, buf + n);
	}
	mm_ctx *raw() { return ctx_; }

private:
	void check(int rc) {
		if (rc != MM_OK) throw std::runtime_error(std::string("mimas_meshgen: ") + mm_last_error_string(ctx_));
	}
	mm_ctx *ctx_ = nullptr;
	std::array<mm_block_entry, MM_TABLE_ENTRIES> table_{};
	bool have_table_ = false;
};

// lib.rs:231-232, the crate's entry point: one mesher per calling thread, as the reference has one
// meshgen thread (client.rs:259).  A mesher failure is fatal in the reference design: it throws.
inline ChunkMesh mesh_for_chunk(const std::array<isize, 3> &offs, const MapChunkData &chunk, const TextureIdCache &cache) {
	thread_local Mesher m(0);
	return m.mesh_for_chunk(offs, chunk, cache);
}
inline void push_block(std::vector<Vertex> &r, const std::array<float, 3> &pos, BlockTextureIds ids, float siz,
                       const std::function<bool(std::array<isize, 3>)> &blocked) {
	thread_local Mesher m(0);
	m.push_block(r, pos, ids, siz, blocked);
}

}  // namespace mimas_meshgen
