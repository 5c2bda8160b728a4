#!/usr/bin/env python
"""BASELINE.json configs[0], [2], [3], [4] on one B200 (configs[1] is bench.py itself; the multi-GPU half of
configs[4] is bench.py under torchrun).  Every number comes with a parity check against the oracle: full
byte compare where the oracle finishes in seconds, otherwise every per-chunk vertex count plus a byte compare
of a random sample of chunks.  Prints one JSON object per config.

    python scripts/run_configs.py [cfg1] [cfg3] [cfg4] [cfg5]      (default: all)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import mimas_b200  # noqa: E402
import oracle  # noqa: E402
from tests.util import load_golden  # noqa: E402

G = load_golden()
OC = oracle.cache_from_entries(G.entries, G.n_known, G.fallback)
CORES = os.cpu_count() or 1
PEAK = 6537.6
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(stream, fn, reps=1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def sample_compare(m, out, chunks_host, offs_host, idx):
    """Byte compare of the chunks `idx` of a device-resident batch result against the oracle."""
    n = out.n_chunks
    ri = m.download(out.d_ranges_i, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    rt = m.download(out.d_ranges_t, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    want = oracle.mesh_batch(chunks_host[idx], offs_host[idx], OC, threads=CORES)
    wi, wt = want["intransparent"].tobytes(), want["transparent"].tobytes()
    for k, c in enumerate(idx):
        ni, nt = int(ri[c]["vertex_count"]), int(rt[c]["vertex_count"])
        gi = m.download(out.d_intransparent + 36 * int(ri[c]["first_vertex"]), 36 * ni).tobytes() if ni else b""
        gt = m.download(out.d_transparent + 36 * int(rt[c]["first_vertex"]), 36 * nt).tobytes() if nt else b""
        a, b = 36 * int(want["first_i"][k]), 36 * int(want["first_t"][k])
        if gi != wi[a:a + 36 * int(want["count_i"][k])] or gt != wt[b:b + 36 * int(want["count_t"][k])]:
            return False
    return True


def checkerboard(n):
    idx = np.arange(4096)
    base = (((idx >> 8) + ((idx >> 4) & 15) + (idx & 15)) & 1).astype(np.uint8)
    par = (np.arange(n) & 1).astype(np.uint8)
    chunks = np.where((base[None, :] ^ par[:, None]) == 1, 5, 0).astype(np.uint8)
    i = np.arange(n)
    offs = np.stack([16 * (i % 32 - 16), 16 * ((i // 32) % 32 - 16), 16 * (i // 1024)], axis=1).astype(np.int64)
    return chunks, offs


def cfg1(m, stream):
    """configs[0]: the spawn chunk (48,32,16) = btchn of (60,40,20) and its surface neighbour (48,32,0),
    one mesh_for_chunk call each (host in, host out) -- latency, against the CPU mesher on one core."""
    res = []
    for cz in (1, 0):
        chunks, offs = mimas_b200.gen_terrain_host(78, (3, 2, cz), (1, 1, 1))
        want = oracle.mesh_batch(chunks, offs, OC)
        got = m.mesh_for_chunk(offs[0], chunks[0])
        ok = (got.intransparent.tobytes() == want["intransparent"].tobytes()
              and got.transparent.tobytes() == want["transparent"].tobytes())
        lat = []
        for _ in range(300):
            t0 = time.perf_counter()
            m.mesh_for_chunk(offs[0], chunks[0])
            lat.append(time.perf_counter() - t0)
        cpu = []
        for _ in range(50):
            cpu.append(oracle.mesh_batch(chunks, offs, OC, threads=1, discard=True)["seconds"])
        res.append({"offs": [int(v) for v in offs[0]], "vertices": int(len(got.intransparent) + len(got.transparent)),
                    "bit_exact": bool(ok), "gpu_call_us_median": 1e6 * float(np.median(lat)),
                    "cpu_1core_us_median": 1e6 * float(np.median(cpu))})
    return {"config": "single 16^3 chunk, synthetic mapgen seed 78 (real mimas mapgen: parity unpinned, DESIGN.md)",
            "note": "one chunk is launch/PCIe-latency bound, not against the roofline", "chunks": res}


def cfg3(m, stream):
    """configs[2]: 16384 solid/air checkerboard chunks (12288 quads each), one call, 43.5 GB of vertices."""
    n = 16384
    chunks, offs = checkerboard(n)
    dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
    m.upload(dc, chunks)
    m.upload(do, offs)
    m.reserve(n, n * 73728, 0)
    out = m.mesh_batch_device(dc, do, n)
    ms = [m.mesh_batch_device(dc, do, n).kernel_ms for _ in range(3)]
    out = m.mesh_batch_device(dc, do, n)
    ri = m.download(out.d_ranges_i, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    counts_ok = bool((ri["vertex_count"] == 73728).all() and out.n_intransparent == n * 73728 and out.n_transparent == 0
                     and (ri["first_vertex"] == np.arange(n, dtype=np.uint64) * 73728).all())
    idx = np.sort(np.random.RandomState(3).choice(n, 24, replace=False))
    sample_ok = sample_compare(m, out, chunks, offs, idx)
    alg = n * (4096 + 24 + 16) + 36 * (out.n_intransparent + out.n_transparent) + 2048 * out.launches // 2
    t = float(np.median(ms)) * 1e-3
    m.dev_free(dc)
    m.dev_free(do)
    return {"config": "16384 worst-case solid/air checkerboard chunks, device resident, one mm_mesh_batch call",
            "chunks_per_s": n / t, "ms": t * 1e3, "vertex_bytes": 36 * int(out.n_intransparent),
            "algorithmic_gbs": alg / t / 1e9, "frac_of_measured_peak": alg / t / 1e9 / PEAK, "launches": int(out.launches),
            "parity": {"all_counts_and_offsets": counts_ok, "sampled_chunks_bit_exact": bool(sample_ok), "sample": len(idx)}}


def cfg4(m, stream):
    """configs[3]: view radius 8/16/32 (z radius 2, config.rs:39) full mesh, then incremental remesh after
    single-block edits: each edit remeshes exactly its chunk (the reference's semantics), in batches of 1/16/256."""
    res = []
    rng = np.random.RandomState(44)
    for r in (8, 16, 32):
        dims = (2 * r + 1, 2 * r + 1, 5)
        n = dims[0] * dims[1] * dims[2]
        dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
        m.gen_terrain_device(78, (-r, -r, -2), dims, dc, do)
        m.sync()
        m.mesh_batch_device(dc, do, n)
        ms = [m.mesh_batch_device(dc, do, n).kernel_ms for _ in range(5)]
        out = m.mesh_batch_device(dc, do, n)
        chunks = m.download(dc, n * 4096).reshape(n, 4096)
        offs = m.download(do, n * 24).view(np.int64).reshape(n, 3)
        full_ok = True
        if n <= 6000:  # full byte compare where the oracle finishes in seconds
            got = m.copy_out(out)
            want = oracle.mesh_batch(chunks, offs, OC, threads=CORES)
            full_ok = (got.intransparent.tobytes() == want["intransparent"].tobytes()
                       and got.transparent.tobytes() == want["transparent"].tobytes())
        else:
            idx = np.sort(rng.choice(n, 512, replace=False))
            full_ok = sample_compare(m, out, chunks, offs, idx)
        t = float(np.median(ms)) * 1e-3
        alg = n * 4136 + 36 * (out.n_intransparent + out.n_transparent) + 2048 * out.launches // 2
        entry = {"radius": r, "chunks": n, "full_mesh_ms": t * 1e3, "chunks_per_s": n / t, "algorithmic_gbs": alg / t / 1e9,
                 "full_mesh_bit_exact": bool(full_ok), "edits": []}
        # K = 256 edits: dig -> air if there is a block, else place stone
        k = 256
        ec = rng.choice(n, k, replace=False)
        cells = rng.randint(0, 4096, size=k)
        edited = chunks[ec].copy()
        for j in range(k):
            edited[j, cells[j]] = 0 if edited[j, cells[j]] != 0 else 5
        eo = offs[ec].copy()
        want = oracle.mesh_batch(edited, eo, OC, threads=CORES)
        for bs in (1, 16, 256):
            lat, ok = [], True
            for s in list(range(0, k, bs)) * (5 if bs == 256 else 1):
                ec_, eo_ = np.ascontiguousarray(edited[s:s + bs]), np.ascontiguousarray(eo[s:s + bs])
                t0 = time.perf_counter()  # the C ABI call itself: host chunks in, host vertex vectors out
                hm = m.mesh_batch_host_raw(ec_.ctypes.data, eo_.ctypes.data, bs)
                lat.append(time.perf_counter() - t0)
                if s < 64 or bs == 256:
                    b = m.mesh_batch(eo_, ec_)
                    a0, a1 = int(want["first_i"][s]), int(want["first_i"][s + bs - 1]) + int(want["count_i"][s + bs - 1])
                    ok &= b.intransparent.tobytes() == want["intransparent"][a0:a1].tobytes() and hm.n_intransparent == a1 - a0
            cpu = oracle.mesh_batch(edited[:bs], eo[:bs], OC, threads=min(bs, CORES), discard=True)["seconds"]
            entry["edits"].append({"batch": bs, "gpu_call_us_median": 1e6 * float(np.median(lat)), "bit_exact": bool(ok),
                                   "cpu_us": 1e6 * cpu, "cpu_threads": min(bs, CORES)})
        res.append(entry)
        m.dev_free(dc)
        m.dev_free(do)
    return {"config": "view-radius sweep 8/16/32 (z radius 2) full mesh, device resident; then 256 single-block edits "
                      "remeshed through the host API in batches of 1/16/256", "sweep": res}


def cfg5(m, stream):
    """configs[4] on one GPU: 1,048,576 terrain chunks (128x128x64 region, 4 GiB of voxels generated on the
    device), one mm_mesh_batch call = 256 launch pairs chained through the device-side carry, output resident."""
    dims = (128, 128, 64)
    n = dims[0] * dims[1] * dims[2]
    dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
    m.gen_terrain_device(78, (-64, -64, -32), dims, dc, do)
    m.sync()
    out = m.mesh_batch_device(dc, do, n)  # sizes the arenas (probe + rerun)
    ms = [m.mesh_batch_device(dc, do, n).kernel_ms for _ in range(2)]
    out = m.mesh_batch_device(dc, do, n)
    t = float(np.median(ms)) * 1e-3
    ri = m.download(out.d_ranges_i, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    rt = m.download(out.d_ranges_t, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    packed = bool((ri["first_vertex"][1:] == np.cumsum(ri["vertex_count"][:-1], dtype=np.uint64)).all()
                  and (rt["first_vertex"][1:] == np.cumsum(rt["vertex_count"][:-1], dtype=np.uint64)).all())
    chunks = m.download(dc, n * 4096).reshape(n, 4096)
    offs = m.download(do, n * 24).view(np.int64).reshape(n, 3)
    # checksum of checksums: total vertex count of all 1M chunks from the CPU mesher (all host threads)
    cpu = oracle.mesh_batch(chunks, offs, OC, threads=CORES, discard=True)
    total_ok = cpu["vertices"] == int(out.n_intransparent + out.n_transparent)
    rng = np.random.RandomState(55)
    idx = np.sort(rng.choice(n, 65536, replace=False))
    ci, ct = oracle.count_batch(chunks[idx], offs[idx], OC)
    counts_ok = bool((ci == ri["vertex_count"][idx]).all() and (ct == rt["vertex_count"][idx]).all())
    sample_ok = sample_compare(m, out, chunks, offs, np.sort(rng.choice(n, 2048, replace=False)))
    alg = n * 4136 + 36 * (out.n_intransparent + out.n_transparent) + 2048 * out.launches // 2
    m.dev_free(dc)
    m.dev_free(do)
    return {"config": "1,048,576 synthetic mapgen chunks (128x128x64 region) on ONE GPU, device resident, one call",
            "chunks_per_s": n / t, "ms": t * 1e3, "vertices": int(out.n_intransparent + out.n_transparent),
            "algorithmic_gbs": alg / t / 1e9, "frac_of_measured_peak": alg / t / 1e9 / PEAK, "launches": int(out.launches),
            "cpu_baseline": {"chunks_per_s": n / cpu["seconds"], "cores": CORES, "kind": "port"},
            "parity": {"streams_tightly_packed_in_chunk_order": packed, "total_vertices_equal_cpu": bool(total_ok),
                       "65536_sampled_counts_equal": counts_ok, "2048_sampled_chunks_bit_exact": bool(sample_ok)}}


def cfg_draw(m, stream):
    """one frame over the view-radius-32 region (21125 chunks): culling (client.rs:679-696)
    + vertex shader (vertex-shader.glsl:18-27) on the device; then the vertex shader over the whole stream."""
    from mimas_b200.render import SHADED_DTYPE, Camera
    r = 32
    dims = (2 * r + 1, 2 * r + 1, 5)
    n = dims[0] * dims[1] * dims[2]
    dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
    m.gen_terrain_device(78, (-r, -r, -2), dims, dc, do)
    out = m.mesh_batch_device(dc, do, n)
    offs = m.download(do, n * 24).view(np.int64).reshape(n, 3)
    frames = []
    for cam in (Camera(pos=(8.0, 8.0, 24.0), pitch=-10.0, yaw=30.0, viewing_range=128.0),
                Camera(pos=(8.0, 8.0, 24.0), pitch=-25.0, yaw=-100.0, viewing_range=512.0),
                Camera(pos=(-200.0, 150.0, 40.0), pitch=-30.0, yaw=20.0, viewing_range=1e6)):
        m.draw_batch(cam)
        ms = [m.draw_batch(cam).kernel_ms for _ in range(7)]
        d = m.draw_batch(cam)
        want = oracle.cull_chunks(offs, cam.get_matrix(), cam.get_perspective(), cam.pos, cam.viewing_range)
        ok = m.download(d.d_draw_list, 4 * d.n_drawn).view(np.uint32).tolist() == want.tolist()
        nv = int(d.n_shaded_i + d.n_shaded_t)
        t = float(np.median(ms)) * 1e-3
        alg = n * (24 + 32) + nv * (36 + 48)  # chunk origins + both range records read; per drawn vertex 36 B in, 48 B out
        frames.append({"viewing_range": cam.viewing_range, "drawn_chunks": int(d.n_drawn), "of": n, "shaded_vertices": nv,
                       "frame_ms": t * 1e3, "algorithmic_gbs": alg / t / 1e9, "draw_list_equals_restatement": bool(ok)})
    # vertex shader alone over every vertex of the region
    nv = int(out.n_intransparent)
    d_out = m.dev_alloc(48 * nv)
    cam = Camera(pos=(1, 2, 3), pitch=5, yaw=6)
    m.set_stream(stream.cuda_stream)
    m.vertex_stage(out.d_intransparent, nv, cam, d_out)
    t = timed(stream, lambda: m.vertex_stage(out.d_intransparent, nv, cam, d_out), reps=20) * 1e-3
    m.set_stream(0)
    idx = np.random.RandomState(2).choice(nv, 4096, replace=False)
    vin = m.download(out.d_intransparent, 36 * nv).view(mimas_b200.meshgen.VERTEX_DTYPE)
    got = m.download(d_out, 48 * nv).view(SHADED_DTYPE)
    want = oracle.vertex_stage(vin, cam.get_matrix(), cam.get_perspective())
    res = {"config": "headless frame over the radius-32 region (21125 terrain chunks): device culling + vertex shader",
           "frames": frames,
           "vertex_stage_all": {"vertices": nv, "ms": t * 1e3, "algorithmic_gbs": nv * 84 / t / 1e9,
                                "frac_of_measured_peak": nv * 84 / t / 1e9 / PEAK, "bit_exact_vs_restatement": bool(got.tobytes() == want.tobytes())}}
    m.dev_free(d_out)
    m.dev_free(dc)
    m.dev_free(do)
    return res


def cfg_halo(m, stream):
    """Halo extension on the bench region (64x32x2 chunks at (-32,-16,-1), seed 78): slabs gathered on the device,
    then the region meshed with and without them; streamed like bench.py (8 batches queued per measurement)."""
    from tests.util import halos_for_box
    dims, c0 = (64, 32, 2), (-32, -16, -1)
    n = dims[0] * dims[1] * dims[2]
    dc, do, dh = m.dev_alloc(n * 4096), m.dev_alloc(n * 24), m.dev_alloc(n * 1536)
    m.gen_terrain_device(78, c0, dims, dc, do)
    m.sync()
    chunks = m.download(dc, n * 4096).reshape(n, 4096)
    offs = m.download(do, n * 24).view(np.int64).reshape(n, 3)
    m.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        m.gather_halos(dc, dims, dh, 0)
        t_gather = timed(stream, lambda: m.gather_halos(dc, dims, dh, 0), reps=20) * 1e-3
        gather_ok = m.download(dh, n * 1536).tobytes() == halos_for_box(chunks, dims, 0).tobytes()
        res = {}
        for name, h in (("reference_mode", None), ("halo_mode", dh)):
            m.mesh_batch_async(dc, do, n, d_halo=h)
            out = m.batch_finish()
            for _ in range(3):
                timed(stream, lambda: m.mesh_batch_async(dc, do, n, d_halo=h), reps=8)
            t = float(np.median([timed(stream, lambda: m.mesh_batch_async(dc, do, n, d_halo=h), reps=8) for _ in range(9)])) * 1e-3
            out = m.batch_finish()
            halo_host = None if h is None else m.download(dh, n * 1536).reshape(n, 6, 256)
            want = oracle.mesh_batch(chunks, offs, OC, threads=CORES, halo=halo_host)
            got = m.copy_out(out)
            ok = (got.intransparent.tobytes() == want["intransparent"].tobytes()
                  and got.transparent.tobytes() == want["transparent"].tobytes())
            nv = int(out.n_intransparent + out.n_transparent)
            alg = n * (4136 + (1536 if h else 0)) + 36 * nv + 2048
            res[name] = {"ms_per_batch": t * 1e3, "chunks_per_s": n / t, "vertices": nv, "algorithmic_gbs": alg / t / 1e9,
                         "frac_of_measured_peak": alg / t / 1e9 / PEAK, "bit_exact_vs_oracle": bool(ok)}
    m.set_stream(0)
    for p in (dc, do, dh):
        m.dev_free(p)
    return {"config": "halo extension: 4096 synthetic mapgen chunks (bench region), slabs gathered on the device",
            "gather_ms": t_gather * 1e3, "gather_equals_numpy": bool(gather_ok), **res}


def cfg_frame(m, stream):
    """SURVEY.md 8f rank 3, optional part: whole headless frames (cull + vertex shader + rasteriser + fragment
    shader) over the radius-32 region at 1920x1080, synthetic textures; one frame is written as a PNG."""
    from mimas_b200.render import Camera, synthetic_textures, write_png
    r = 32
    dims = (2 * r + 1, 2 * r + 1, 5)
    n = dims[0] * dims[1] * dims[2]
    dc, do = m.dev_alloc(n * 4096), m.dev_alloc(n * 24)
    m.gen_terrain_device(78, (-r, -r, -2), dims, dc, do)
    m.mesh_batch_device(dc, do, n)
    m.set_textures(synthetic_textures(80, crossed=range(30, 44)))
    W, H = 1920, 1080
    frames = []
    for k, cam in enumerate((Camera(pos=(8.0, 8.0, 24.0), pitch=10.0, yaw=30.0, viewing_range=128.0, aspect_ratio=W / H),
                             Camera(pos=(8.0, 8.0, 40.0), pitch=35.0, yaw=-100.0, viewing_range=512.0, aspect_ratio=W / H),
                             Camera(pos=(-200.0, 150.0, 60.0), pitch=30.0, yaw=20.0, viewing_range=1e6, aspect_ratio=W / H))):
        d = m.draw_batch(cam)
        nv = int(d.n_shaded_i + d.n_shaded_t)
        img, _ = m.raster_frame(d.d_shaded or 0, nv, W, H, 40.0, 60.0 if k else 1e9, want_ms=True)
        ms = [m.raster_frame(d.d_shaded or 0, nv, W, H, 40.0, 60.0, want_ms=True)[1] for _ in range(5)]
        dms = [m.draw_batch(cam).kernel_ms for _ in range(5)]
        drawn = float(((img >> 24) & 0xFF > 0).mean())
        frames.append({"viewing_range": cam.viewing_range, "drawn_chunks": int(d.n_drawn), "triangles": nv // 3,
                       "cull_and_vertex_ms": float(np.median(dms)), "raster_ms": float(np.median(ms)),
                       "covered_fraction": drawn})
        if k == 0:
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            write_png(os.path.join(ROOT, "gpurun_out", "frame.png"), img[::2, ::2])
    # parity of one small frame against the oracle chain is tests/test_raster.py; here only a checksum of self-consistency
    m.dev_free(dc)
    m.dev_free(do)
    return {"config": "headless frames 1920x1080 over the radius-32 region (21125 terrain chunks), synthetic textures",
            "frames": frames}


def main():
    which = [a for a in sys.argv[1:] if a.startswith("cfg")] or ["cfg1", "cfg3", "cfg4", "cfg5", "cfg_draw", "cfg_halo", "cfg_frame"]
    m = mimas_b200.Mesher(0, mimas_b200.default_cache())
    stream = torch.cuda.Stream()
    fns = {"cfg1": cfg1, "cfg3": cfg3, "cfg4": cfg4, "cfg5": cfg5, "cfg_draw": cfg_draw, "cfg_halo": cfg_halo, "cfg_frame": cfg_frame}
    for w in which:
        t0 = time.time()
        r = fns[w](m, stream)
        r["name"] = w
        r["wall_s"] = round(time.time() - t0, 1)
        print(json.dumps(r), flush=True)
    m.close()


if __name__ == "__main__":
    main()
