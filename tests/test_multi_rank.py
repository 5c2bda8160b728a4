"""The N > 1 path on CPU: one process per rank, `gloo`, world size 2.  Chunks are independent, so ranks mesh
disjoint shards with no collective on the data path; torch.distributed only carries the barrier and the
max / sum reductions of the bench, and the reference arm runs on rank 0 alone."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import bench
    import mimas_b200
    import oracle
    from tests.util import load_golden, oracle_cache
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    barrier, rmax, rsum = bench.make_reducers(world, "cpu")
    # the shard of this rank: the bench's rotating batches, shrunk to 4x4x4 chunks each
    origins = [bench.region_origin(rank, r) for r in range(3)]
    verts = 0
    for o in origins:
        chunks, offs = mimas_b200.gen_terrain_host(bench.SEED, o, (4, 4, 4))
        verts += oracle.mesh_batch(chunks, offs, oracle_cache(load_golden()), threads=2, discard=True)["vertices"]
    barrier()
    total = rsum(float(verts))
    slowest = rmax(float(rank + 1))
    gathered = [None] * world
    dist.all_gather_object(gathered, origins)
    q.put((rank, verts, total, slowest, gathered))
    dist.destroy_process_group()


def test_two_ranks_mesh_disjoint_shards_and_reduce_over_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, v0, t0, s0, g0), (r1, v1, t1, s1, g1) = res
    assert (r0, r1) == (0, 1)
    assert t0 == t1 == float(v0 + v1) and v0 > 0 and v1 > 0       # whole-job aggregate = sum over ranks
    assert s0 == s1 == 2.0                                         # timing = max over ranks
    tiles = [tuple(o) for o in g0[0] + g0[1]]
    assert g0 == g1 and len(set(tiles)) == len(tiles)              # shards are disjoint 32x32-chunk tiles
    xs = sorted(set(o[1] for o in tiles))
    assert xs[1] - xs[0] >= 32


def test_reference_arm_under_torchrun_prints_one_line_from_rank0():
    """bench.py --impl reference launched like the driver launches it for N > 1: rank 0 alone runs the CPU mesher
    and prints the JSON line, the other rank exits 0 without work."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
           "--steps", "1", "--warmup", "1"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["gpu_launches"] == 0
    assert d["metric"] == "chunks_meshed_per_sec" and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0


def test_million_shards_partition_the_region_and_every_workload_has_a_config():
    """configs[4] under N ranks: the 128x128x64 region is cut along y into N disjoint boxes that add up to 1,048,576
    chunks (strong scaling); and every bench workload names its configuration."""
    sys.path.insert(0, ROOT)
    import bench
    for world in (1, 2, 4, 8):
        seen, total = set(), 0
        for rank in range(world):
            c0, dims = bench.million_shard(rank, world)
            assert dims[0] == 128 and dims[2] == 64 and dims[1] == 128 // world and c0[0] == -64 and c0[2] == -32
            ys = set(range(c0[1], c0[1] + dims[1]))
            assert not (ys & seen)
            seen |= ys
            total += dims[0] * dims[1] * dims[2]
        assert total == 1 << 20 and seen == set(range(-64, 64))
    for wl in bench.WORKLOADS:
        cfg = bench.workload_config(wl, 4096, 2)
        assert cfg["workload"] and cfg["l2"]
    c, o = bench.checkerboard(5, parity0=1)
    assert c.shape == (5, 4096) and set(np.unique(c)) == {0, 5} and (c[0] != c[1]).all() and o.shape == (5, 3)
