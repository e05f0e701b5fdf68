"""Developer diagnostic (GPU box): large (PCG) path vs oracle."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from oracle import ba_cpu  # noqa: E402
from visualodometry_b200 import BAConfig, BundleAdjuster, synth  # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)) if a.size else 0.0


def K_of(p):
    return np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1]])


def check(name, p, iters=8, pcg_max=200, cpu=True):
    print(f"=== {name}: nc={p.nc} np={p.np_} no={p.no}", flush=True)
    cfg = BAConfig(max_iters=iters, solver="pcg", record_history=True, pcg_max_iters=pcg_max)
    ba = BundleAdjuster(K_of(p), cfg)
    t0 = time.time()
    ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    print(f"  set_problem {time.time() - t0:.3f}s")
    if cpu and p.nc <= 300:
        ix = ba_cpu.index(p)
        for k, ref in (("PM_PERM", ix.pm_perm), ("PT_PTR", ix.pt_ptr), ("BLK_PTR", ix.blk_ptr), ("BLK_COL", ix.blk_col)):
            print(f"  index {k}: equal={bool(np.array_equal(ba.dump(k, ref.shape, np.int32), ref))}")
        st = ba_cpu.stage(p, mu=1e4, solver="pcg", pcg_max_iters=pcg_max)
        n = st["S"].shape[0]
        for k, ok, shp in (("S", "S", (n, n)), ("GRED", "gred", (n,)), ("DC", "dc", (n,))):
            print(f"  stage {k}: rel_err={rel(ba.dump(k, shp), st[ok]):.3e}")
        print("  scalars", ba.dump("SCALARS", (4,)), "oracle cost", st["cost"], "trial", st["trial_cost"])
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for rep in range(3):
        ba.reset(); torch.cuda.synchronize()
        ev[0].record(); ba.solve(wait=False); ev[1].record(); torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    sm = ba.summaries()[0]
    poses, pts = ba.get()
    print(f"  engine: {sm}  ms {ts} -> {p.no * sm.iterations / min(ts) * 1e3:.4e} obs*iter/s")
    print("  cost hist", ba.dump("HIST_COST", (1, 64))[0, :iters])
    print("  rmse", synth.rmse_px(p), "->", synth.rmse_px(p, poses, pts))
    if cpu:
        t0 = time.time()
        po, xo, so = ba_cpu.solve(p, max_iters=iters, solver="pcg", pcg_max_iters=pcg_max, threads=0)
        dt = time.time() - t0
        print(f"  oracle: it={so.iterations} acc={so.accepted} c1={so.final_cost:.9e} pcg={so.pcg_iters_total} "
              f"{dt:.2f}s -> {p.no * so.iterations / dt:.4e} obs*iter/s")
        print("  cost hist rel", rel(ba.dump("HIST_COST", (1, 64))[0, :iters], np.array(so.cost_hist[:iters])),
              "poses abs", float(np.abs(poses - po).max()), "pts abs", float(np.abs(pts - xo).max()))
    ba.close()


if __name__ == "__main__":
    check("random-small", synth.make_large(2, n_cams=40, n_pts=2000, k=5, mode="random", path_len=40.0))
    check("local-small", synth.make_large(2, n_cams=40, n_pts=2000, k=5, mode="local", path_len=40.0))
    check("random-mid", synth.make_large(3, n_cams=200, n_pts=20000, k=5, mode="random", path_len=200.0))
    check("C5-local", synth.make_large(0, mode="local"), iters=20, cpu=True)
    check("C5-random", synth.make_large(0, mode="random"), iters=20, cpu=False)
