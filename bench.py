#!/usr/bin/env python
"""bench.py -- chunks meshed per second on B200, against the HBM roofline and the CPU mesher.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload terrain|checkerboard]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One step = one pass of the mesher (an analyse + an expand kernel launch) over one batch of 4096 synthetic mapgen
chunks (BASELINE.json configs[1]).  Every rank owns one GPU and meshes its own batches (weak
scaling, no collective on the data path); torch.distributed is used only for the barrier and the
max-over-ranks of the timings.  Prints ONE JSON line on rank 0.
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
L2_BYTES = 126e6


def region_origin(rank: int, r: int):
    """Chunk coordinates of the r-th rotating batch of a rank: disjoint 32x32 tiles of the world."""
    return (-16 + 32 * r, -16 + 32 * rank, -2)


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
    idx = np.arange(4096)
    base = ((idx >> 8) + ((idx >> 4) & 15) + (idx & 15)) & 1
    chunks = np.empty((n, 4096), dtype=np.uint8)
    offs = np.empty((n, 3), dtype=np.int64)
    for i in range(n):
        chunks[i] = np.where((base ^ ((i + parity0) & 1)) == 1, 5, 0)
        offs[i] = (16 * (i % 32 - 16), 16 * ((i // 32) % 32 - 16), 16 * (i // 1024))
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
    """DRAM bytes per launch of the mesher kernel from the committed ncu capture, if any."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            pass
    return None


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
                                          "-lms", "50", "-i", str(index)], stdout=subprocess.PIPE,
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


def cpu_mesher_rate(chunks, offs, threads, min_seconds=3.0, max_reps=40):
    """The oracle (C restatement of mimas-meshgen/lib.rs:231-446) on `threads` host threads, building
    every mesh in Vec-like growing buffers and discarding it.  Returns (chunks/s, reps)."""
    import oracle
    from tests.util import load_golden
    g = load_golden()
    cache = oracle.cache_from_entries(g.entries, g.n_known, g.fallback)
    oracle.mesh_batch(chunks[:256], offs[:256], cache, threads=threads, discard=True)  # warm
    total, reps = 0.0, 0
    while total < min_seconds and reps < max_reps:
        total += oracle.mesh_batch(chunks, offs, cache, threads=threads, discard=True)["seconds"]
        reps += 1
    return len(chunks) * reps / total, reps


def run_reference(args, rank, world):
    """--impl reference: the reference mesher's CPU implementation (C restatement; the Rust original
    cannot be built here: no rustc/cargo) on all host threads, same workload, K steps."""
    if rank != 0:
        return
    import mimas_b200
    import oracle
    from tests.util import load_golden
    g = load_golden()
    cache = oracle.cache_from_entries(g.entries, g.n_known, g.fallback)
    cores = os.cpu_count() or 1
    chunks, offs = make_host_batch(args.workload, 0, 0)
    for _ in range(max(args.warmup, 1)):
        oracle.mesh_batch(chunks, offs, cache, threads=cores, discard=True)
    t = 0.0
    for _ in range(args.steps):
        t += oracle.mesh_batch(chunks, offs, cache, threads=cores, discard=True)["seconds"]
    value = len(chunks) * args.steps / t
    sample = f"{len(chunks)} chunks per step ({args.workload}), every mesh built and dropped, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8+f32", "data": "synthetic",
        "config": workload_config(args.workload, len(chunks)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(workload, n):
    if workload == "terrain":
        return {"workload": f"batch of {n} synthetic mapgen terrain chunks (32x32x4 region, seed {SEED}), "
                            "no halo (reference semantics), default game block table",
                "l2": f"{ROTATE} distinct input batches rotated ({ROTATE * n * 4096 / 1e6:.0f} MB > L2); "
                      "each step also writes a fresh >126 MB vertex stream"}
    return {"workload": f"batch of {n} worst-case solid/air checkerboard chunks (12288 quads each)",
            "l2": "each step writes a vertex stream far larger than L2"}


def make_host_batch(workload, rank, r):
    import mimas_b200
    if workload == "terrain":
        return mimas_b200.gen_terrain_host(SEED, region_origin(rank, r), REGION)
    return checkerboard(1024, parity0=r)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="terrain", choices=["terrain", "checkerboard"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="host-in/host-out steps (default: min(steps, 20))")
    ap.add_argument("--no-verify", action="store_true", help="skip the pre-timing parity check")
    ap.add_argument("--quick", action="store_true", help="tuning runs: device-resident timing only (no e2e, no CPU baseline)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

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

    barrier, max_over_ranks, sum_over_ranks = make_reducers(world, "cuda")

    mesher = mimas_b200.Mesher(local_rank, mimas_b200.default_cache())
    stream = torch.cuda.Stream()
    mesher.set_stream(stream.cuda_stream)
    sm_count, ctas_per_sm, threads = mesher.device_info()

    # ---- inputs: ROTATE distinct batches, generated on the device, resident in HBM ----------
    n = N_CHUNKS if args.workload == "terrain" else 1024
    rot = ROTATE if args.workload == "terrain" else 2
    d_chunks = [torch.empty(n * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
    d_offs = [torch.empty(n * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
    for r in range(rot):
        if args.workload == "terrain":
            mesher.gen_terrain_device(SEED, region_origin(rank, r), REGION, d_chunks[r].data_ptr(), d_offs[r].data_ptr())
        else:
            c, o = checkerboard(n, parity0=r)
            d_chunks[r].copy_(torch.from_numpy(c.reshape(-1)))
            d_offs[r].copy_(torch.from_numpy(o.reshape(-1)))
    mesher.sync()
    torch.cuda.synchronize()

    # ---- correctness gate + algorithmic bytes (oracle as the checker, before any timing) ------
    import oracle
    from tests.util import load_golden
    g = load_golden()
    ocache = oracle.cache_from_entries(g.entries, g.n_known, g.fallback)
    verts, iso_ms = [], []
    for r in range(rot):
        out = mesher.mesh_batch_device(d_chunks[r].data_ptr(), d_offs[r].data_ptr(), n)
        verts.append(out.n_intransparent + out.n_transparent)
        iso_ms.append(out.kernel_ms)  # one batch alone on the GPU, events around its two kernels
        if not args.no_verify and (r == 0 or args.workload == "terrain" and rank == 0 and r < 3):
            got = mesher.copy_out(out)
            hc = d_chunks[r].cpu().numpy().reshape(n, 4096)
            ho = d_offs[r].cpu().numpy().reshape(n, 3)
            want = oracle.mesh_batch(hc, ho, ocache, threads=os.cpu_count() or 1)
            ok = (got.intransparent.tobytes() == want["intransparent"].tobytes()
                  and got.transparent.tobytes() == want["transparent"].tobytes()
                  and list(got.ranges_i["vertex_count"]) == list(want["count_i"]))
            if not ok:
                raise SystemExit(f"PARITY FAILURE on batch {r}: CUDA output differs from the oracle; no number reported")
    # 4096 (voxels) + 24 (offs) + 36 * vertices + 16 (two range records) per chunk,
    # + the 2 KiB block table once per launch
    alg_bytes = [n * (4096 + 24 + 16) + 36 * v + 2048 for v in verts]

    # ---- device-resident timed region ----------------------------------------------------------
    launches0 = mesher.launch_count()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    windows = []
    with torch.cuda.stream(stream):
        for i in range(args.warmup):
            mesher.mesh_batch_async(d_chunks[i % rot].data_ptr(), d_offs[i % rot].data_ptr(), n)
        mesher.batch_finish()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches0 = mesher.launch_count()
        barrier()
        torch.cuda.synchronize()
        t_host0 = time.time()
        ev0.record(stream)
        for i in range(args.steps):
            mesher.mesh_batch_async(d_chunks[i % rot].data_ptr(), d_offs[i % rot].data_ptr(), n)
        ev1.record(stream)
        torch.cuda.synchronize()
        barrier()
        windows.append((t_host0, time.time()))
        last = mesher.batch_finish()
    gpu_launches = mesher.launch_count() - launches0
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    ms_per_step = ms_total / args.steps
    value = world * n * args.steps / (ms_total * 1e-3)
    step_bytes = float(np.mean([alg_bytes[i % rot] for i in range(args.steps)]))
    achieved = step_bytes / (ms_per_step * 1e-3) / 1e9
    vert_bytes = float(np.mean([36 * verts[i % rot] for i in range(args.steps)]))

    # ---- end to end through the C ABI with host buffers (H2D + kernel + D2H every step) -----------
    e2e_steps = 1 if args.quick else (args.e2e_steps or min(args.steps, 20))
    hc, ho = make_host_batch(args.workload, rank, 0)
    pin_c, pin_o = mesher.host_alloc(n * 4096), mesher.host_alloc(n * 24)
    pin_c[:] = hc.reshape(-1)
    pin_o[:] = ho.reshape(-1).view(np.uint8)
    mesher.set_stream(0)  # the library's own stream for the synchronous host API
    for _ in range(3):
        hm = mesher.mesh_batch_host_raw(pin_c.ctypes.data, pin_o.ctypes.data, n)
    d2h = 36 * (hm.n_intransparent + hm.n_transparent) + 2 * 16 * n + 16
    h2d = n * (4096 + 24)
    barrier()
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(e2e_steps):
        hm = mesher.mesh_batch_host_raw(pin_c.ctypes.data, pin_o.ctypes.data, n)
    t1 = time.time()
    windows.append((t0, t1))
    e2e_s = max_over_ranks(t1 - t0)
    e2e_value = world * n * e2e_steps / e2e_s
    clocks = sampler.stop(windows) if sampler else None

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only) ---------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.quick:
        cores = os.cpu_count() or 1
        rate_all, reps = cpu_mesher_rate(hc, ho, cores)
        rate_1, _ = cpu_mesher_rate(hc[:1024], ho[:1024], 1, min_seconds=1.0)
        cpu = {"value": rate_all, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{reps} x the same {n}-chunk batch, every mesh built and dropped "
                         "(C restatement of mimas-meshgen; the Rust original is unavailable on this box)",
               "single_thread_value": rate_1}

    if rank == 0:
        peak, peak_src = peaks()
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8+f32", "data": "synthetic", "config": workload_config(args.workload, n),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(), "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": step_bytes,
                         "kernel": "mm::analyse_kernel + mm::expand_kernel (one pair per step; consecutive "
                                   "steps overlap by programmatic dependent launch)",
                         "isolated_batch_ms": float(np.median(iso_ms[1:] or iso_ms)),
                         "vertex_out_gbs": vert_bytes / (ms_per_step * 1e-3) / 1e9,
                         "frac_of_nominal_8000": achieved / 8000.0},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                    "api": "mm_mesh_batch_host (pinned host chunks in, pinned host vertex vectors out)"},
            "gpu_launches": int(gpu_launches),
            "clocks": clocks,
            "grid": {"sms": sm_count, "ctas_per_sm": ctas_per_sm, "threads_per_cta": threads},
            "vertices_per_step": float(np.mean(verts)),
        }
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
