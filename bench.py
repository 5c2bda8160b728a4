#!/usr/bin/env python
"""bench.py -- chunks meshed per second on B200, against the HBM roofline and the CPU mesher.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload W]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workloads (BASELINE.json `configs`, one JSON line each, same schema):
    terrain            configs[1] (default): 4096 synthetic mapgen chunks per step, streamed through the output ring;
                       under torchrun every rank meshes its own batches (weak scaling)
    single             configs[0]: ONE chunk per step (the spawn chunk and the one below it) -- latency, not roofline
    checkerboard16384  configs[2]: 16384 worst-case solid/air checkerboard chunks per step (43.5 GB of vertices)
    radius8|16|32      configs[3]: full mesh of a view-radius region (1445 / 5445 / 21125 chunks) per step, one call
    edits              configs[3]: remesh after single-block edits (each edit remeshes exactly its chunk)
    million            configs[4]: 1,048,576 mapgen chunks (128x128x64 region) per step; under torchrun the region is
                       cut into N contiguous shards (STRONG scaling, no collective)
    mapgen             like terrain, but the chunks come from the device restatement of mimas-server's mapgen phase one
                       (mm_gen_mapgen_device; parity of that generator to the reference is unpinned) instead of the look-alike
A step = one pass of the mesher over the workload's batch, inputs resident in HBM; `e2e` = the same through the
host API (host chunks in, H2D, kernels, D2H, host vertex vectors out).  Every workload is first compared with
the CPU oracle (parity gate) -- no number is printed for a mesh that differs.  torch.distributed carries only
the barrier and the max/sum of the timings.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "chunks_meshed_per_sec"
UNIT = "chunks/s"
SEED = 78                      # mimas-common/config.rs:37 default mapgen seed
REGION = (32, 32, 4)           # 4096 chunks: 32 x 32 columns, z-chunks -2..+1
N_CHUNKS = REGION[0] * REGION[1] * REGION[2]
ROTATE = 12                    # distinct input batches cycled through: 12 x 16 MiB > 126 MB of L2
RING = int(os.environ.get("MM_BENCH_RING", "3"))  # output slots of the streamed workload: three batches stay retrievable
L2_BYTES = 126e6
WORKLOADS = ["terrain", "single", "checkerboard16384", "radius8", "radius16", "radius32", "edits", "million", "checkerboard", "mapgen"]
MILLION = (128, 128, 64)       # configs[4]: chunk region centred on the origin


def region_origin(rank: int, r: int):
    """Chunk coordinates of the r-th rotating batch of a rank: disjoint 32x32 tiles of the world."""
    return (-16 + 32 * r, -16 + 32 * rank, -2)


def million_shard(rank: int, world: int):
    """configs[4] under N ranks: the 128x128x64 region cut along y into N contiguous boxes (every box holds all
    z layers, so the shards cost the same).  Returns (origin in chunks, dims)."""
    ny = MILLION[1] // world
    return (-MILLION[0] // 2, -MILLION[1] // 2 + rank * ny, -MILLION[2] // 2), (MILLION[0], ny, MILLION[2])


def make_reducers(world: int, device: str):
    """barrier / max-over-ranks / sum-over-ranks: the only uses of torch.distributed in the bench (the data
    path has no collective: chunks are independent, every rank meshes its own shard)."""
    import torch
    import torch.distributed as dist

    def barrier():
        if world > 1:
            dist.barrier()

    def reduce(x: float, op) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=op)
        return float(t.item())

    return barrier, (lambda x: reduce(x, dist.ReduceOp.MAX)), (lambda x: reduce(x, dist.ReduceOp.SUM))


def checkerboard(n, parity0=0):
    """(x+y+z+parity)&1 ? stone : air, parity alternating per chunk, origins on a 32x32xk grid."""
    idx = np.arange(4096)
    base = (((idx >> 8) + ((idx >> 4) & 15) + (idx & 15)) & 1).astype(np.uint8)
    par = ((np.arange(n) + parity0) & 1).astype(np.uint8)
    chunks = np.where((base[None, :] ^ par[:, None]) == 1, 5, 0).astype(np.uint8)
    i = np.arange(n)
    offs = np.stack([16 * (i % 32 - 16), 16 * ((i // 32) % 32 - 16), 16 * (i // 1024)], axis=1).astype(np.int64)
    return chunks, offs


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """Steady-state DRAM bytes per step of the mesher kernels from the committed ncu capture, if any
    (profiles/roofline_traffic.json, written by scripts/ncu_steady.py)."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return d.get("dram_bytes_per_step"), None
        except Exception:
            pass
    return None, None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed regions (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, windows):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ts, line in self.samples:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            if not any(a - 0.05 <= ts <= b + 0.05 for a, b in windows):
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- the CPU mesher (oracle: C restatement of mimas-meshgen/lib.rs:231-446) --------------------------------

def oracle_cache():
    import oracle
    from tests.util import load_golden
    g = load_golden()
    return oracle.cache_from_entries(g.entries, g.n_known, g.fallback)


def cpu_mesher_rate(chunks, offs, threads, min_seconds=3.0, max_reps=0):
    """`threads` host threads sharing one atomic work counter, every mesh built in Vec-like growing buffers and
    dropped.  Returns (chunks/s, reps)."""
    import oracle
    cache = oracle_cache()
    oracle.mesh_batch(chunks[:256], offs[:256], cache, threads=threads, discard=True)  # warm
    total, reps = 0.0, 0
    max_reps = max_reps or (40 if len(chunks) >= 1024 else 4000)
    while total < min_seconds and reps < max_reps:
        total += oracle.mesh_batch(chunks, offs, cache, threads=threads, discard=True)["seconds"]
        reps += 1
    return len(chunks) * reps / total, reps


def host_sample(workload, rank=0, world=1):
    """Host-side chunks of a workload for the CPU arms and the e2e leg: the whole batch where it is small, a
    bounded sample of it otherwise (the second value says what)."""
    import mimas_b200
    if workload == "terrain":
        return mimas_b200.gen_terrain_host(SEED, region_origin(rank, 0), REGION) + ("the whole 4096-chunk batch",)
    if workload == "mapgen":   # the restated mapgen on the host (oracle side: CPU arms only)
        from mimas_b200 import mapgen as pm
        from oracle import mapgen as om
        c, o, _ = om.gen_region(SEED, region_origin(rank, 0), REGION, pm.to_c(pm.default_mapgen_params()))
        return c, o, "the whole 4096-chunk batch"
    if workload == "single":
        a = mimas_b200.gen_terrain_host(SEED, (3, 2, 1), (1, 1, 1))
        b = mimas_b200.gen_terrain_host(SEED, (3, 2, 0), (1, 1, 1))
        return np.concatenate([a[0], b[0]]), np.concatenate([a[1], b[1]]), "the two chunks, one call each"
    if workload in ("checkerboard16384", "checkerboard"):
        return checkerboard(1024) + ("1024 of the 16384 checkerboard chunks",)
    if workload.startswith("radius") or workload == "edits":
        r = 8 if workload == "edits" else int(workload[6:])
        c, o = mimas_b200.gen_terrain_host(SEED, (-r, -r, -2), (2 * r + 1, 2 * r + 1, 5))
        if workload == "edits":
            c, o, _ = apply_edits(c, o)
            return c, o, "the 256 edited chunks of the radius-8 region"
        return c, o, f"the whole region ({len(c)} chunks)"
    if workload == "million":
        c0, dims = million_shard(rank, world)
        sub = (dims[0], min(dims[1], 8), 16)     # 128 x 8 x 16 = 16384 chunks around the surface
        return mimas_b200.gen_terrain_host(SEED, (c0[0], c0[1], -8), sub) + (f"{sub[0] * sub[1] * sub[2]} chunks of the region (z-chunks -8..7)",)
    raise ValueError(workload)


def apply_edits(chunks, offs, k=256, seed=44):
    """configs[3]: k single-block edits on distinct chunks -- dig (-> air) where there is a block, place stone where
    there is none (mimas-client/client.rs:946-950, 1023-1024).  Returns the edited chunks, their origins, indices."""
    rng = np.random.RandomState(seed)
    ec = rng.choice(len(chunks), k, replace=False)
    cells = rng.randint(0, 4096, size=k)
    edited = chunks[ec].copy()
    for j in range(k):
        edited[j, cells[j]] = 0 if edited[j, cells[j]] != 0 else 5
    return edited, offs[ec].copy(), ec


def workload_config(workload, n, world=1):
    base = {"terrain": f"batch of {n} synthetic mapgen terrain chunks (32x32x4 region, seed {SEED}), "
                       "no halo (reference semantics), default game block table",
            "mapgen": f"batch of {n} chunks of the restated mimas-server mapgen phase one (32x32x4 region, seed {SEED}; "
                      "generated on the device, parity of the generator unpinned), no halo, default game block table",
            "single": "single 16^3 chunk per call: synthetic mapgen chunk at offs (48,32,16) and (48,32,0), alternating",
            "checkerboard16384": "batch of 16384 worst-case solid/air checkerboard chunks (12288 quads each), one call",
            "checkerboard": f"batch of {n} worst-case solid/air checkerboard chunks (12288 quads each)",
            "edits": "radius-8 region resident; 256 single-block edits, every edited chunk remeshed in full (one batch of 256)",
            "million": f"1,048,576 synthetic mapgen chunks (128x128x64 region, seed {SEED}) cut along y into {world} "
                       f"contiguous shard(s) of {n} chunks, one call per shard, no collective"}
    if workload.startswith("radius"):
        r = int(workload[6:])
        base[workload] = f"view radius {r} (z radius 2): full mesh of {n} synthetic mapgen chunks in one call"
    l2 = {"mapgen": f"{ROTATE} distinct input batches rotated ({ROTATE * n * 4096 / 1e6:.0f} MB > L2); "
                    f"each step also writes a fresh >126 MB vertex stream into a ring of {RING} arenas",
          "terrain": f"{ROTATE} distinct input batches rotated ({ROTATE * n * 4096 / 1e6:.0f} MB > L2); "
                     f"each step also writes a fresh >126 MB vertex stream into a ring of {RING} arenas",
          "single": "latency workload: nothing to flush (one 4 KiB chunk per call)",
          "edits": "latency workload: L2 flushed (256 MB written) between timed steps",
          "million": "inputs (4 GiB / N) and outputs (40 GB / N) far larger than L2"}
    return {"workload": base[workload],
            "l2": l2.get(workload, "each step writes a vertex stream far larger than L2" if workload.startswith("checker")
                         else "L2 flushed (256 MB written) between timed steps, each timed alone with CUDA events")}


def run_reference(args, rank, world):
    """--impl reference: the reference mesher's CPU implementation (C restatement; the Rust original cannot be built
    here: no rustc/cargo) on all host threads, on a bounded sample of the same workload, K steps."""
    if rank != 0:
        return
    import oracle
    cache = oracle_cache()
    cores = os.cpu_count() or 1
    chunks, offs, what = host_sample(args.workload, 0, world)
    threads = 1 if args.workload == "single" else cores
    for _ in range(max(args.warmup, 1)):
        oracle.mesh_batch(chunks, offs, cache, threads=threads, discard=True)
    t = 0.0
    for _ in range(args.steps):
        t += oracle.mesh_batch(chunks, offs, cache, threads=threads, discard=True)["seconds"]
    value = len(chunks) * args.steps / t
    sample = f"{what}, per step; every mesh built and dropped, {threads} thread(s) sharing one work counter"
    n_cfg = {"terrain": N_CHUNKS, "million": MILLION[0] * MILLION[1] * MILLION[2] // world}.get(args.workload, len(chunks))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
        "scaling": "strong" if args.workload == "million" else "weak",
        "vs_baseline": None, "dtype": "u8+f32", "data": "synthetic",
        "config": workload_config(args.workload, n_cfg, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---- parity gates --------------------------------------------------------------------------------------------

def full_compare(mesher, out, hc, ho, cache):
    import oracle
    got = mesher.copy_out(out)
    want = oracle.mesh_batch(hc, ho, cache, threads=os.cpu_count() or 1)
    return (got.intransparent.tobytes() == want["intransparent"].tobytes()
            and got.transparent.tobytes() == want["transparent"].tobytes()
            and list(got.ranges_i["vertex_count"]) == list(want["count_i"])
            and list(got.ranges_t["vertex_count"]) == list(want["count_t"]))


def sample_compare(mesher, out, fetch, idx, cache):
    """Byte compare of the chunks `idx` of a device-resident result against the oracle; fetch(idx) -> host chunks, offs."""
    import mimas_b200
    import oracle
    n = out.n_chunks
    ri = mesher.download(out.d_ranges_i, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    rt = mesher.download(out.d_ranges_t, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
    packed = bool((ri["first_vertex"][1:] == np.cumsum(ri["vertex_count"][:-1], dtype=np.uint64)).all()
                  and (rt["first_vertex"][1:] == np.cumsum(rt["vertex_count"][:-1], dtype=np.uint64)).all())
    hc, ho = fetch(idx)
    want = oracle.mesh_batch(hc, ho, cache, threads=os.cpu_count() or 1)
    wi, wt = want["intransparent"].tobytes(), want["transparent"].tobytes()
    ok = packed
    for k, c in enumerate(idx):
        ni, nt = int(ri[c]["vertex_count"]), int(rt[c]["vertex_count"])
        ok = ok and ni == int(want["count_i"][k]) and nt == int(want["count_t"][k])
        if not ok:
            break
        gi = mesher.download(out.d_intransparent + 36 * int(ri[c]["first_vertex"]), 36 * ni).tobytes() if ni else b""
        gt = mesher.download(out.d_transparent + 36 * int(rt[c]["first_vertex"]), 36 * nt).tobytes() if nt else b""
        a, b = 36 * int(want["first_i"][k]), 36 * int(want["first_t"][k])
        ok = ok and gi == wi[a:a + 36 * ni] and gt == wt[b:b + 36 * nt]
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (default: per workload, 300 for terrain)")
    ap.add_argument("--warmup", type=int, default=0)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="terrain", choices=WORKLOADS)
    ap.add_argument("--e2e-steps", type=int, default=0, help="host-in/host-out steps (default: per workload)")
    ap.add_argument("--no-verify", action="store_true", help="skip the pre-timing parity check")
    ap.add_argument("--quick", action="store_true", help="tuning runs: device-resident timing only (no e2e, no CPU baseline)")
    ap.add_argument("--consume", action="store_true",
                    help="terrain: every streamed batch is handed to the headless draw (cull + vertex shader) while the next ones mesh")
    args = ap.parse_args()
    wl = args.workload
    defaults = {"mapgen": (300, 20), "terrain": (300, 20), "single": (400, 20), "checkerboard16384": (5, 3), "checkerboard": (40, 5),
                "radius8": (40, 5), "radius16": (40, 5), "radius32": (30, 5), "edits": (40, 5), "million": (5, 3)}
    if wl == "terrain" and args.consume:
        defaults["terrain"] = (60, 6)
    if args.steps <= 0:
        args.steps = defaults[wl][0] if args.impl == "b200" else 5
    if args.warmup <= 0:
        args.warmup = defaults[wl][1] if args.impl == "b200" else 1
    args.warmup = max(args.warmup, 3 if args.impl == "b200" else 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    import mimas_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the mesher has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        try:  # one core set per rank: the launch cadence of a rank must not depend on its neighbours
            cores = sorted(os.sched_getaffinity(0))
            per = max(1, len(cores) // world)
            os.sched_setaffinity(0, set(cores[local_rank * per:(local_rank + 1) * per]) or set(cores))
        except (AttributeError, OSError):
            pass

    barrier, max_over_ranks, sum_over_ranks = make_reducers(world, "cuda")
    mesher = mimas_b200.Mesher(local_rank, mimas_b200.default_cache())
    stream = torch.cuda.Stream()
    mesher.set_stream(stream.cuda_stream)
    sm_count, ctas_per_sm, threads = mesher.device_info()
    cache = oracle_cache()
    streamed = wl in ("terrain", "checkerboard", "mapgen")   # batches queued back to back through the ring
    strong = wl == "million"

    # ---- inputs, resident in HBM ---------------------------------------------------------------------------
    if wl in ("terrain", "mapgen"):
        n, rot = N_CHUNKS, ROTATE
        boxes = [(region_origin(rank, r), REGION) for r in range(rot)]
    elif wl == "checkerboard":
        n, rot, boxes = 1024, 2, None
    elif wl == "checkerboard16384":
        n, rot, boxes = 16384, 1, None
    elif wl == "single":
        n, rot, boxes = 1, 2, [((3, 2, 1), (1, 1, 1)), ((3, 2, 0), (1, 1, 1))]
    elif wl.startswith("radius") or wl == "edits":
        r = 8 if wl == "edits" else int(wl[6:])
        dims = (2 * r + 1, 2 * r + 1, 5)
        n, rot, boxes = dims[0] * dims[1] * dims[2], 1, [((-r, -r, -2), dims)]
    else:  # million
        c0, dims = million_shard(rank, world)
        n, rot, boxes = dims[0] * dims[1] * dims[2], 1, [(c0, dims)]
    d_chunks = [torch.empty(n * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
    d_offs = [torch.empty(n * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
    for r in range(rot):
        if wl == "mapgen":
            mesher.gen_mapgen_device(SEED, boxes[r][0], boxes[r][1], d_chunks[r].data_ptr(), d_offs[r].data_ptr())
        elif boxes is not None:
            mesher.gen_terrain_device(SEED, boxes[r][0], boxes[r][1], d_chunks[r].data_ptr(), d_offs[r].data_ptr())
        else:
            c, o = checkerboard(n, parity0=r)
            d_chunks[r].copy_(torch.from_numpy(c.reshape(-1)))
            d_offs[r].copy_(torch.from_numpy(o.reshape(-1)))
    mesher.sync()
    torch.cuda.synchronize()
    if wl == "edits":   # the edited chunks become the (device-resident) batch that is remeshed
        hc_all = d_chunks[0].cpu().numpy().reshape(n, 4096)
        ho_all = d_offs[0].cpu().numpy().reshape(n, 3)
        ec, eo, _ = apply_edits(hc_all, ho_all)
        n = len(ec)
        d_chunks = [torch.from_numpy(ec.reshape(-1)).cuda()]
        d_offs = [torch.from_numpy(eo.reshape(-1)).cuda()]

    def fetch(idx, r=0):
        idx = torch.as_tensor(np.asarray(idx), device="cuda", dtype=torch.int64)
        return (d_chunks[r].view(-1, 4096)[idx].cpu().numpy(), d_offs[r].view(-1, 3)[idx].cpu().numpy())

    # ---- correctness gate + algorithmic bytes (oracle as the checker, before any timing) -----------------------
    if streamed:
        mesher.ring_configure(RING)
    verts, iso_ms, launches_per_step, parity = [], [], 0, "skipped (--no-verify)"
    for r in range(rot):
        out = mesher.mesh_batch_device(d_chunks[r].data_ptr(), d_offs[r].data_ptr(), n)   # first call sizes the arenas
        out = mesher.mesh_batch_device(d_chunks[r].data_ptr(), d_offs[r].data_ptr(), n)
        verts.append(out.n_intransparent + out.n_transparent)
        iso_ms.append(out.kernel_ms)  # one batch alone on the GPU, events around its kernels
        launches_per_step = out.launches
        if args.no_verify or not (r == 0 or (wl in ("terrain", "mapgen") and rank == 0 and r < 3) or wl == "single"):
            continue
        if n <= 6000:
            ok = full_compare(mesher, out, d_chunks[r].cpu().numpy().reshape(n, 4096), d_offs[r].cpu().numpy().reshape(n, 3), cache)
            parity = "every byte of both vertex streams and all range records equal the oracle's"
        else:
            k = 2048 if wl == "million" else 512
            idx = np.sort(np.random.RandomState(3 + rank).choice(n, k, replace=False))
            ok = sample_compare(mesher, out, lambda i, r=r: fetch(i, r), idx, cache)
            parity = f"streams tightly packed in chunk order; {k} sampled chunks byte-exact with the oracle"
            if wl == "million" and rank == 0:   # every count of a 65536-chunk sample
                import oracle
                idx = np.sort(np.random.RandomState(9).choice(n, 16384, replace=False))
                hc, ho = fetch(idx)
                ci, ct = oracle.count_batch(hc, ho, cache)
                ri = mesher.download(out.d_ranges_i, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
                rt = mesher.download(out.d_ranges_t, 16 * n).view(mimas_b200.meshgen.RANGE_DTYPE)
                ok = ok and bool((ri["vertex_count"][idx] == ci).all() and (rt["vertex_count"][idx] == ct).all())
                parity += "; 16384 more sampled vertex counts equal"
        if not ok:
            raise SystemExit(f"PARITY FAILURE on batch {r} of workload {wl}: CUDA output differs from the oracle; no number reported")
    # 4096 (voxels) + 24 (offs) + 36 * vertices + 16 (two range records) per chunk,
    # + the 2 KiB block table once per launch pair
    pairs = max(1, (n + 4095) // 4096)
    alg_bytes = [n * (4096 + 24 + 16) + 36 * v + 2048 * pairs for v in verts]

    # ---- device-resident timed region ----------------------------------------------------------------------
    sampler = ClockSampler(local_rank)   # every rank samples its own GPU; rank 0's goes into `clocks`, all of them into `ranks`
    windows = []
    consumed = None
    flush = None if streamed or wl in ("single", "million", "checkerboard16384") else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    with torch.cuda.stream(stream):
        if streamed and not args.consume:
            ptr_c = [t.data_ptr() for t in d_chunks]
            ptr_o = [t.data_ptr() for t in d_offs]
            many = mesher.prepare_many(ptr_c, ptr_o, [n] * rot)   # the argument arrays of the one submission call
            for i in range(args.warmup):
                mesher.mesh_batch_async(ptr_c[i % rot], ptr_o[i % rot], n)
            mesher.batch_finish()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            launches0 = mesher.launch_count()
            barrier()
            torch.cuda.synchronize()
            t_host0 = time.time()
            ev0.record(stream)
            # ONE foreign call enqueues the K batches (round robin over the rotating inputs): the host language is
            # not in the ~35 us launch cadence (mm_mesh_batches_submit)
            first = mesher.submit_many(prepared=many, repeat=args.steps // rot) if args.steps >= rot else None
            for i in range(args.steps - (args.steps // rot) * rot if first else args.steps):
                mesher.mesh_batch_async(ptr_c[i % rot], ptr_o[i % rot], n)
            ev1.record(stream)
            torch.cuda.synchronize()
            barrier()
            windows.append((t_host0, time.time()))
            mesher.batch_finish()
            ms_total = ev0.elapsed_time(ev1)
        elif streamed:
            # every batch is consumed: batch k is culled + vertex-shaded on the consumer stream (mm_draw_ticket)
            # while batches k+1, k+2 are meshed on the mesher's stream
            from mimas_b200.render import Camera
            cam = Camera(pos=(8.0, 8.0, 24.0), pitch=-10.0, yaw=30.0, viewing_range=1e6).to_c()
            tk = [mesher.submit(d_chunks[i % rot].data_ptr(), d_offs[i % rot].data_ptr(), n) for i in range(RING - 1)]
            for i in range(args.warmup):
                tk.append(mesher.submit(d_chunks[(i + RING - 1) % rot].data_ptr(), d_offs[(i + RING - 1) % rot].data_ptr(), n))
                mesher.draw_ticket(tk.pop(0), cam)
            launches0 = mesher.launch_count()
            barrier()
            torch.cuda.synchronize()
            shaded = 0
            t_host0 = time.time()
            for i in range(args.steps):
                # the frame over batch k is enqueued, batch k + RING - 1 is submitted while it runs, then the frame is collected
                mesher.draw_ticket_async(tk.pop(0), cam)
                j = i + args.warmup + RING - 1
                tk.append(mesher.submit(d_chunks[j % rot].data_ptr(), d_offs[j % rot].data_ptr(), n))
                d = mesher.draw_collect()
                shaded += int(d.n_shaded_i + d.n_shaded_t)
            torch.cuda.synchronize()
            t_host1 = time.time()
            barrier()
            windows.append((t_host0, t_host1))
            for t in tk:
                mesher.batch_wait(t)
            ms_total = 1e3 * (t_host1 - t_host0)   # host clock: the consumer's stream is part of the step
            consumed = {"consumer": "mm_draw_ticket: cull + vertex shader over every batch (client.rs:679-696, vertex-shader.glsl)",
                        "shaded_vertices_per_step": shaded / args.steps, "timed_with": "host clock around the loop"}
        else:
            # one synchronous call per step, each timed alone with CUDA events (mm_batch_out.kernel_ms), L2 flushed between
            for i in range(args.warmup):
                mesher.mesh_batch_device(d_chunks[i % rot].data_ptr(), d_offs[i % rot].data_ptr(), n)
            launches0 = mesher.launch_count()
            barrier()
            torch.cuda.synchronize()
            t_host0 = time.time()
            ms_total = 0.0
            for i in range(args.steps):
                if flush is not None:
                    flush.fill_(i & 0xFF)
                    torch.cuda.synchronize()
                ms_total += mesher.mesh_batch_device(d_chunks[i % rot].data_ptr(), d_offs[i % rot].data_ptr(), n).kernel_ms
            torch.cuda.synchronize()
            barrier()
            windows.append((t_host0, time.time()))
    gpu_launches = mesher.launch_count() - launches0
    own_ms_total = ms_total
    ms_total = max_over_ranks(ms_total)
    ms_per_step = ms_total / args.steps
    n_job = world * n   # strong scaling: the shards add up to the whole region; weak: every rank its own batch
    value = n_job * args.steps / (ms_total * 1e-3)
    step_bytes = float(np.mean([alg_bytes[i % rot] for i in range(args.steps)]))
    achieved = step_bytes / (ms_per_step * 1e-3) / 1e9
    vert_bytes = float(np.mean([36 * verts[i % rot] for i in range(args.steps)]))

    # ---- end to end through the C ABI with host buffers (H2D + kernel + D2H every step) ----------------------
    if wl == "mapgen":   # the product's own generator made these chunks on the device: fetch them (the oracle's mapgen is for the CPU arms)
        hc = d_chunks[0].cpu().numpy().reshape(n, 4096)
        ho = d_offs[0].cpu().numpy().reshape(n, 3)
        e2e_what = "the whole 4096-chunk batch"
    else:
        hc, ho, e2e_what = host_sample(wl, rank, world)
    ne = len(hc)
    e2e_steps = 1 if args.quick else (args.e2e_steps or {"terrain": min(args.steps, 20), "single": 200, "million": 3}.get(wl, 5))
    pin_c, pin_o = mesher.host_alloc(ne * 4096), mesher.host_alloc(ne * 24)
    pin_c[:] = hc.reshape(-1)
    pin_o[:] = ho.reshape(-1).view(np.uint8)
    mesher.set_stream(0)  # the library's own stream for the synchronous host API
    per_call = 1 if wl == "single" else ne
    calls = ne // per_call

    def host_step():
        tot, path = 0, 0
        for k in range(calls):
            hm = mesher.mesh_batch_host_raw(pin_c.ctypes.data + k * per_call * 4096, pin_o.ctypes.data + k * per_call * 24, per_call)
            tot += hm.n_intransparent + hm.n_transparent
            path = hm.path
        return tot, path
    for _ in range(3):
        e2e_verts, e2e_path = host_step()
    d2h = 36 * e2e_verts + 2 * 16 * ne + 16
    h2d = ne * (4096 + 24)
    barrier()
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(e2e_steps):
        host_step()
    t1 = time.time()
    windows.append((t0, t1))
    e2e_s = max_over_ranks(t1 - t0)
    e2e_value = world * ne * e2e_steps / e2e_s
    clocks = sampler.stop(windows) if sampler else None
    per_rank = None
    if world > 1:   # outside every timed region: which GPU was the slowest, and was it throttled (the value follows the slowest rank)
        import torch.distributed as dist
        per_rank = [None] * world
        dist.all_gather_object(per_rank, {"rank": rank, "ms_per_step": own_ms_total / args.steps, "clocks": clocks})

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only): a bounded sample of the workload -----------
    cpu = None
    if rank == 0 and world == 1 and not args.quick:
        cores = os.cpu_count() or 1
        use = 1 if wl == "single" else cores
        rate_all, reps = cpu_mesher_rate(hc, ho, use)
        rate_1, _ = cpu_mesher_rate(hc[:1024], ho[:1024], 1, min_seconds=1.0)
        cpu = {"value": rate_all, "unit": UNIT, "cores": use, "kind": "port",
               "sample": f"{reps} x {e2e_what}, every mesh built and dropped "
                         "(C restatement of mimas-meshgen; the Rust original is unavailable on this box)",
               "single_thread_value": rate_1}

    if rank == 0:
        peak, peak_src = peaks()
        traffic, ncu_gbs = ncu_traffic() if wl == "terrain" and not args.consume else (None, None)
        if traffic:   # ncu-counted steady-state DRAM bytes per step over the step time measured here, un-profiled
            ncu_gbs = traffic / (ms_per_step * 1e-3) / 1e9
        latency = wl in ("single", "edits")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "u8+f32", "data": "synthetic", "config": workload_config(wl, n, world),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "ncu_dram_gbs": ncu_gbs, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": step_bytes,
                         "kernel": "mm::analyse_kernel + mm::expand_kernel (one pair per 4096 chunks; consecutive "
                                   "pairs overlap by programmatic dependent launch)",
                         "isolated_batch_ms": float(np.median(iso_ms[1:] or iso_ms)),
                         "vertex_out_gbs": vert_bytes / (ms_per_step * 1e-3) / 1e9,
                         "frac_of_nominal_8000": achieved / 8000.0},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps, "chunks_per_step": ne,
                    "sample": e2e_what, "path": mimas_b200.meshgen.HOST_PATH_NAMES.get(e2e_path, str(e2e_path)),
                    "api": ("mm_mesh_for_chunk-shaped calls, one chunk each" if wl == "single" else "mm_mesh_batch_host")
                           + " (pinned host chunks in, pinned host vertex vectors out)"},
            "gpu_launches": int(gpu_launches),
            "clocks": clocks,
            "grid": {"sms": sm_count, "ctas_per_sm": ctas_per_sm, "threads_per_cta": threads},
            "vertices_per_step": float(np.mean(verts)),
            "parity": parity,
            "launches_per_step": int(launches_per_step),
        }
        if latency:
            line["roofline"]["note"] = "latency-bound workload (launch + PCIe round trip), reported for completeness, not graded against the roofline"
            line["latency_us"] = {"device_resident_call": 1e3 * ms_per_step, "host_call": 1e6 * e2e_s / e2e_steps / calls}
        if per_rank:
            line["ranks"] = per_rank
        if consumed:
            line["consumed"] = consumed
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    mesher.host_free(pin_c)
    mesher.host_free(pin_o)
    mesher.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
