#!/usr/bin/env python
"""Device timeline of consecutive overlapped steps (run on the GPU box): where a step's time goes.
    python scripts/timeline.py [steps]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/timeline.py [steps]
Under torchrun every rank traces its own GPU during the same window (after a barrier) and prints one summary
line: is a longer step at N = 8 a gap before the analysis launch (host cadence) or longer kernels?"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import mimas_b200  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 24
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
m = mimas_b200.Mesher(local, mimas_b200.default_cache())
n, rot = 4096, 6
dc = [torch.empty(n * 4096, dtype=torch.uint8, device="cuda") for _ in range(rot)]
do = [torch.empty(n * 3, dtype=torch.int64, device="cuda") for _ in range(rot)]
for r in range(rot):
    m.gen_terrain_device(78, (-16 + 32 * r, -16 + 32 * rank, -2), (32, 32, 4), dc[r].data_ptr(), do[r].data_ptr())
m.sync()
m.ring_configure(3)
for r in range(3):
    m.mesh_batch_device(dc[r].data_ptr(), do[r].data_ptr(), n)  # sizes the arenas of every slot
pc, po = [t.data_ptr() for t in dc], [t.data_ptr() for t in do]
m.submit_many(pc, po, [n] * rot, repeat=4)
m.batch_finish()
m.debug_trace(steps)
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
m.submit_many(pc, po, [n] * rot, repeat=max(1, steps // rot))
m.batch_finish()
t = m.debug_trace(0).astype(np.int64)
t0 = t[0, 0, 0]
a, e = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
if world == 1:
    print("step |  A.start  A.workdone  A.end |  E.start  E.pastwait  E.end | period(us)")
    for k in range(len(t)):
        per = "" if k == 0 else f"{e[k, 3] - e[k - 1, 3]:7.1f}"
        print(f"{k:4d} | {a[k, 0]:8.1f} {a[k, 2]:10.1f} {a[k, 3]:7.1f} | {e[k, 0]:8.1f} {e[k, 1]:10.1f} {e[k, 3]:7.1f} | {per}")
k0 = 3
period = np.diff(e[k0:, 3])
summary = {"rank": rank, "steps": int(len(t)), "period_us_mean": float(period.mean()), "period_us_max": float(period.max()),
           "analysis_us_mean": float((a[k0:, 2] - a[k0:, 0]).mean()), "expansion_us_mean": float((e[k0:, 3] - e[k0:, 1]).mean()),
           # analysis k+1 may start once expansion k is past its wait: the lag is the launch side (host cadence + hand-over)
           "analysis_start_after_expand_pastwait_us_mean": float((a[k0 + 1:, 0] - e[k0:-1, 1]).mean()),
           "expand_pastwait_after_analysis_end_us_mean": float((e[k0:, 1] - a[k0:, 3]).mean())}
print(json.dumps(summary), flush=True)
m.close()
if world > 1:
    dist.destroy_process_group()
