"""CPU-only checks of the boundary and the host logic: the C-ABI library loads and exports every
symbol include/mimas_meshgen_cuda.h declares (no compute without a GPU), the product-side block-table
derivation agrees with the oracle's and with the committed fixture, the synthetic terrain generator is
deterministic."""
import ctypes as C
import hashlib
import os
import re

import numpy as np
import pytest

from tests.util import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = load_golden()
REF_TOML = "/root/reference/mimas-server/game-params.toml"


def header_symbols():
    text = open(os.path.join(ROOT, "include", "mimas_meshgen_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mm_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from mimas_b200._lib import SYMBOLS, lib
    l = lib()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(l, s), f"{s} declared in the header but not exported"
    assert sorted(SYMBOLS) == syms
    assert l.mm_abi_version() == 1


def test_struct_layouts_match_header():
    from mimas_b200 import meshgen as mg
    assert mg.VERTEX_DTYPE.itemsize == 36 and mg.VERTEX_DTYPE.fields["tex_pos"][1] == 4
    assert mg.VERTEX_DTYPE.fields["position"][1] == 12 and mg.VERTEX_DTYPE.fields["normal"][1] == 24
    assert mg.RANGE_DTYPE.itemsize == 16 and mg.BLOCK_ENTRY_DTYPE.itemsize == 12 and mg.CUBE_DTYPE.itemsize == 24
    assert C.sizeof(mg._BatchOut) == 64 and C.sizeof(mg._HostMesh) == 56


def test_no_device_means_error_not_fallback():
    """On a machine without a GPU the context cannot be created; nothing silently runs on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import mimas_b200
    from mimas_b200._lib import MMError
    with pytest.raises(MMError):
        mimas_b200.Mesher(0)


def test_table_from_cache_fills_unknown_ids_with_fallback():
    import mimas_b200
    tic = mimas_b200.TextureIdCache.from_table_rows(G.entries, G.fallback)
    t = tic.to_table()
    assert len(t) == 256
    for i, (kind, s, tp, b, x) in enumerate(G.entries):
        assert (int(t[i]["kind"]), int(t[i]["id_sides"]), int(t[i]["id_top"]), int(t[i]["id_bottom"]),
                int(t[i]["id_crossed"])) == (kind, s, tp, b, x)
    for i in range(len(G.entries), 256):
        assert int(t[i]["kind"]) == 1
        assert (int(t[i]["id_sides"]), int(t[i]["id_top"]), int(t[i]["id_bottom"])) == G.fallback
    assert tic.get_bl_tex_ids(200) == tic.fallback_id and tic.get_mesh_draw_style(200) is None
    assert tic.get_bl_tex_ids(42) is None and tic.get_mesh_draw_style(42).id == 30


@pytest.mark.skipif(not os.path.exists(REF_TOML), reason="reference checkout not present (GPU box)")
def test_product_table_derivation_matches_fixture_and_oracle():
    """TextureIdCache::from_hdl (lib.rs:87-104) with the headless id allocator (assets.rs:140-201)."""
    import mimas_b200
    from oracle import block_table as obt
    text = open(REF_TOML).read()
    params = mimas_b200.load_params_from_toml(text)
    tic = mimas_b200.TextureIdCache.from_hdl(params, mimas_b200.HeadlessAssets())
    t = tic.to_table()
    rows = [(int(e["kind"]), int(e["id_sides"]), int(e["id_top"]), int(e["id_bottom"]), int(e["id_crossed"]))
            for e in t[:len(params.block_params)]]
    assert rows == G.entries
    assert (tic.fallback_id.id_sides, tic.fallback_id.id_top, tic.fallback_id.id_bottom) == G.fallback
    assert params.name_id_map.id_to_name == G.table["names"]
    assert tic.inv_texture_ids == G.table["inv_texture_ids"]
    assert rows == [tuple(r) for r in obt.default_cache_from_toml(text).to_entries()[:len(rows)]]


def test_default_table_shipped_with_the_package():
    import json
    shipped = json.load(open(os.path.join(ROOT, "mimas_b200", "data", "default_block_table.json")))
    assert [tuple(r) for r in shipped["entries"]] == G.entries and tuple(shipped["fallback"]) == G.fallback


def test_terrain_generator_is_deterministic_and_plausible():
    import mimas_b200
    a, oa = mimas_b200.gen_terrain_host(78, (-4, -4, -2), (8, 8, 4))
    b, ob = mimas_b200.gen_terrain_host(78, (-4, -4, -2), (8, 8, 4))
    assert a.tobytes() == b.tobytes() and oa.tobytes() == ob.tobytes()
    assert hashlib.sha256(a.tobytes()).hexdigest() == open(
        os.path.join(ROOT, "tests", "golden", "terrain_8x8x4_seed78.sha256")).read().strip()
    assert oa[0].tolist() == [-64, -64, -32] and oa[1].tolist() == [-48, -64, -32]
    # a sub-region is the same bytes as the region generated on its own
    c, oc = mimas_b200.gen_terrain_host(78, (-4, -4, -1), (8, 8, 1))
    assert c.tobytes() == a[64:128].tobytes() and oc.tobytes() == oa[64:128].tobytes()
    ids = set(np.unique(a))
    assert {0, 1, 5} <= ids and ids <= set(range(56))


def test_header_compiles_as_c_and_struct_sizes_match_the_bindings(tmp_path):
    """include/mimas_meshgen_cuda.h is plain C (the cgo / Rust-FFI side needs nothing else), and the ctypes
    mirrors of its structs have the sizes the C compiler gives them."""
    import subprocess
    from mimas_b200 import meshgen as mg
    from mimas_b200 import render as rd
    names = ["mm_vertex", "mm_range", "mm_block_entry", "mm_cube", "mm_batch_out", "mm_host_mesh", "mm_camera",
             "mm_shaded_vertex", "mm_draw_out", "mm_frame"]
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "mimas_meshgen_cuda.h"\nint main(void) {\n'
                   + "".join(f'\tprintf("{n} %zu\\n", sizeof({n}));\n' for n in names) + "\treturn 0;\n}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True)
    got = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    want = {"mm_vertex": mg.VERTEX_DTYPE.itemsize, "mm_range": mg.RANGE_DTYPE.itemsize,
            "mm_block_entry": mg.BLOCK_ENTRY_DTYPE.itemsize, "mm_cube": mg.CUBE_DTYPE.itemsize,
            "mm_batch_out": C.sizeof(mg._BatchOut), "mm_host_mesh": C.sizeof(mg._HostMesh),
            "mm_camera": C.sizeof(rd.CCamera), "mm_shaded_vertex": rd.SHADED_DTYPE.itemsize,
            "mm_draw_out": C.sizeof(rd.DrawOut), "mm_frame": C.sizeof(rd.CFrame)}
    assert {k: int(v) for k, v in got.items()} == want
