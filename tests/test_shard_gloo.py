"""N>1 host logic on CPU: the landmark partition used for the observation-sharded C5 path, checked
with a world_size-2 gloo all-reduce.  Each rank linearises ONLY its landmark range (oracle, the
checker); the all-reduced reduced camera system must equal the unsharded one (SURVEY §8 row e:
S = B - sum_p ... is linear in the landmark ranges once the damping is taken out, mu -> inf)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from visualodometry_b200 import shard, synth

MU_INF = 1e300


def test_ranges_tile_and_balance():
    p = synth.make_large(4, n_cams=60, n_pts=5000, k=5, mode="random", path_len=60.0)
    # uneven observation counts: drop a third of the observations of the first half of the landmarks
    keep = ~((p.obs_pt < 2500) & (np.arange(p.no) % 3 == 0))
    op = p.obs_pt[keep]
    for n in (1, 2, 3, 8):
        r = shard.point_ranges(op, p.np_, n)
        assert r[0][0] == 0 and r[-1][1] == p.np_ and all(a[1] == b[0] for a, b in zip(r, r[1:]))
        cnt = [int(((op >= lo) & (op < hi)).sum()) for lo, hi in r]
        assert max(cnt) - min(cnt) <= 8          # balanced by observations, up to one landmark's worth
    assert shard.window_ranges(4096, 8) == [(512 * i, 512 * (i + 1)) for i in range(8)]
    assert [hi - lo for lo, hi in shard.window_ranges(10, 4)] == [3, 3, 2, 2]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import ba_cpu
    p = synth.make_large(5, n_cams=24, n_pts=600, k=5, mode="random", path_len=24.0)
    lo, hi = shard.point_ranges(p.obs_pt, p.np_, world)[rank]
    sub = shard.shard_problem(p, lo, hi)
    st = ba_cpu.stage(sub, mu=MU_INF)
    # undamped pieces are additive over landmark ranges: S (incl. the B_c diagonal blocks), g_red, cost
    buf = torch.from_numpy(np.concatenate([st["S"].ravel(), st["gred"], [st["cost"]]]))
    dist.all_reduce(buf)
    if rank == 0:
        full = ba_cpu.stage(p, mu=MU_INF)
        ref = np.concatenate([full["S"].ravel(), full["gred"], [full["cost"]]])
        out.put(float(np.abs(buf.numpy() - ref).max() / np.abs(ref).max()))
    dist.destroy_process_group()


def test_sharded_reduced_system_equals_unsharded_gloo(oracle):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for q in procs:
        q.start()
    for q in procs:
        q.join(120)
        assert q.exitcode == 0
    assert out.get() < 1e-12
