#!/usr/bin/env python
"""Steady-state DRAM traffic of the streamed terrain workload (bench.py's default), for `roofline.traffic`.

On the GPU box, three runs of this script:
  plain:   python scripts/ncu_steady.py                       (must exit 0 first; also prints the CUDA-event time per step)
  kernels: ncu --cache-control none --clock-control none --metrics $M -s <warm launches> -c 24 --csv --log-file ... python scripts/ncu_steady.py
           every launch of 12 chained steps, L2 NOT flushed between kernels: dirty lines a kernel leaves in L2 are
           written back during the next kernels and counted there, so the per-step SUM is the steady-state traffic
  range:   ncu --replay-mode range --cache-control none --clock-control none --metrics $M --csv --log-file ... python scripts/ncu_steady.py
           the 12 steps between cudaProfilerStart/Stop as ONE measured range, kernels overlapping as they do in the bench
with M = dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum,gpu__time_duration.sum
Here (no GPU): python scripts/ncu_steady.py --summarise gpurun_out/<kernels.csv> gpurun_out/<range.csv> -> profiles/roofline_traffic.json
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
STEPS, WARM, N = 12, 24, 4096


def run():
    import torch
    import bench
    import mimas_b200
    m = mimas_b200.Mesher(0, mimas_b200.default_cache())
    st = torch.cuda.Stream()
    m.set_stream(st.cuda_stream)
    rot = bench.ROTATE
    dc = [torch.empty(N * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
    do = [torch.empty(N * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
    for r in range(rot):
        m.gen_terrain_device(bench.SEED, bench.region_origin(0, r), bench.REGION, dc[r].data_ptr(), do[r].data_ptr())
    m.sync()
    m.ring_configure(bench.RING)
    verts = []
    for r in range(rot):
        m.mesh_batch_device(dc[r].data_ptr(), do[r].data_ptr(), N)
        out = m.mesh_batch_device(dc[r].data_ptr(), do[r].data_ptr(), N)
        verts.append(out.n_intransparent + out.n_transparent)
    many = m.prepare_many([t.data_ptr() for t in dc], [t.data_ptr() for t in do], [N] * rot)
    with torch.cuda.stream(st):
        m.submit_many(prepared=many, repeat=WARM // rot)
        m.batch_finish()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.profiler.start()
        e0.record(st)
        m.submit_many(prepared=many, repeat=STEPS // rot)
        e1.record(st)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        m.batch_finish()
    alg = sum(N * (4096 + 24 + 16) + 36 * verts[i % rot] + 2048 for i in range(STEPS)) / STEPS
    print(json.dumps({"steps": STEPS, "event_us_per_step": 1e3 * e0.elapsed_time(e1) / STEPS, "algorithmic_bytes_per_step": alg,
                      "launches_before_the_measured_steps": 4 * rot + 2 * WARM + rot}))  # terrain gen + 2 x 2 per gate call + warm-up pairs
    m.close()


def read_ncu_csv(path):
    rows = list(csv.reader(l for l in open(path, errors="replace") if l.startswith('"')))
    hdr = rows[0]
    out = []
    for r in rows[1:]:
        d = dict(zip(hdr, r))
        out.append(d)
    return out


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "sector": 32, "usecond": 1, "us": 1, "nsecond": 1e-3, "ns": 1e-3,
                                        "msecond": 1e3, "ms": 1e3, "second": 1e6, "s": 1e6}.get(unit, 1)


def summarise(kernels_csv, range_csv, plain_json=None):
    res = {"how": "ncu --cache-control none, streamed terrain workload (scripts/ncu_steady.py); see the docstring"}
    if plain_json and os.path.exists(plain_json):
        res["plain_run"] = json.loads(open(plain_json).read().strip().splitlines()[-1])
    if kernels_csv and os.path.exists(kernels_csv):
        rows = read_ncu_csv(kernels_csv)
        per = {}
        for r in rows:
            k = (r["ID"], r["Kernel Name"])
            per.setdefault(k, {})[r["Metric Name"]] = to_bytes(r["Metric Value"], r["Metric Unit"])
        launches = [dict(kernel=k[1], **v) for k, v in per.items()]
        mesher = [l for l in launches if "analyse_kernel" in l["kernel"] or "expand_kernel" in l["kernel"]]
        steps = len(mesher) // 2
        tot = lambda key: sum(l.get(key, 0.0) for l in mesher)
        res["per_kernel_capture"] = {
            "launches": len(mesher), "steps": steps,
            "dram_read_bytes_per_step": tot("dram__bytes_read.sum") / steps, "dram_write_bytes_per_step": tot("dram__bytes_write.sum") / steps,
            "l2_write_bytes_per_step": tot("lts__t_sectors_op_write.sum") / steps,
            "serialised_kernel_us_per_step": tot("gpu__time_duration.sum") / steps,
            "note": "kernels serialised by ncu (no overlap): the traffic is steady state, the time is not the bench's"}
    if range_csv and os.path.exists(range_csv):
        rows = read_ncu_csv(range_csv)
        d = {r["Metric Name"]: to_bytes(r["Metric Value"], r["Metric Unit"]) for r in rows}
        res["range_capture"] = {"steps": STEPS, "dram_read_bytes_per_step": d.get("dram__bytes_read.sum", 0) / STEPS,
                                "dram_write_bytes_per_step": d.get("dram__bytes_write.sum", 0) / STEPS,
                                "l2_write_bytes_per_step": d.get("lts__t_sectors_op_write.sum", 0) / STEPS,
                                "range_us_per_step": d.get("gpu__time_duration.sum", 0) / STEPS}
    src = res.get("range_capture") or res.get("per_kernel_capture")
    if src:
        res["dram_bytes_per_step"] = src["dram_read_bytes_per_step"] + src["dram_write_bytes_per_step"]
        res["note"] = ("bench.py divides dram_bytes_per_step by the step time it measures itself (CUDA events, un-profiled) "
                       "to report roofline.ncu_dram_gbs")
    print(json.dumps(res, indent=1))
    return res


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--summarise":
        r = summarise(*(sys.argv[2:5]))
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json"), "w") as f:
            json.dump(r, f, indent=1)
    else:
        run()
