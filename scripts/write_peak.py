#!/usr/bin/env python
"""Pure-write and copy bandwidth of this GPU with plain torch ops (context for the roofline: the mesher's
traffic is 90 % writes, MEASURED_PEAKS.json's figure is a copy).  Prints one JSON line."""
import json
import torch

n = 1 << 30
a = torch.empty(n, dtype=torch.bfloat16, device="cuda")
b = torch.empty(n, dtype=torch.bfloat16, device="cuda")


def best(fn, reps=10):
    t = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        t.append(e0.elapsed_time(e1) * 1e-3)
    return min(t)


for _ in range(3):
    a.fill_(1.0); b.copy_(a)
torch.cuda.synchronize()
fill = 2 * n / best(lambda: a.fill_(2.0)) / 1e9
zero = 2 * n / best(lambda: a.zero_()) / 1e9
copy = 4 * n / best(lambda: b.copy_(a)) / 1e9
read = 2 * n / best(lambda: a.sum()) / 1e9
print(json.dumps({"fill_write_only_gbs": fill, "zero_write_only_gbs": zero, "copy_read_plus_write_gbs": copy, "sum_read_only_gbs": read}))
