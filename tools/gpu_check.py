"""Developer diagnostic (GPU box): engine vs oracle, stage by stage, printing every difference.
Not a test (tests/ hold the asserting versions); used to localise bugs in one gpurun call."""
import json
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
    d = np.abs(a - b).max() if a.size else 0.0
    s = max(np.abs(b).max() if b.size else 0.0, 1e-300)
    return float(d / s)


def K_of(p):
    return np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1]])


def check_problem(name, p, loss="trivial", iters=10):
    print(f"=== {name}: nc={p.nc} np={p.np_} no={p.no} loss={loss}", flush=True)
    cfg = BAConfig(max_iters=iters, loss=loss, record_history=True)
    ba = BundleAdjuster(K_of(p), cfg)
    ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    ix = ba_cpu.index(p)
    for k, ref in (("PM_PERM", ix.pm_perm), ("PT_PTR", ix.pt_ptr), ("CM_PERM", ix.cm_perm),
                   ("CAM_PTR", ix.cam_ptr), ("FIDX", ix.fidx)):
        got = ba.dump(k, ref.shape, np.int32)
        print(f"  index {k}: equal={bool((got == ref).all())}")
    st = ba_cpu.stage(p, mu=cfg.init_radius, loss=loss)
    n = st["S"].shape[0]
    shapes = dict(RESIDUALS=("r", (p.no, 2)), WEIGHTS=("w", (p.no,)), JC=("Jc", (p.no, 2, 6)),
                  JP=("Jp", (p.no, 2, 3)), B=("B", (p.nc, 6, 6)), GC=("gc", (p.nc, 6)), C=("C", (p.np_, 6)),
                  GP=("gp", (p.np_, 3)), S=("S", (n, n)), GRED=("gred", (n,)), DC=("dc", (n,)),
                  DP=("dp", (p.np_, 3)), SCALARS=("scalars", (4,)))
    for k, (ok, shp) in shapes.items():
        try:
            got = ba.dump(k, shp)
            ref = st[ok]
            if k in ("B", "GC"):   # fixed cameras: engine leaves B/gc zero
                free = p.fixed_cam == 0
                got, ref = got[free], ref[free]
            print(f"  stage {k}: rel_err={rel(got, ref):.3e}")
            if k == "SCALARS":
                print("     engine", got, "oracle", ref)
        except Exception as e:  # noqa: BLE001
            print(f"  stage {k}: FAILED {e}")
    t0 = time.time()
    sm = ba.solve()[0]
    poses, pts = ba.get()
    t1 = time.time()
    po, xo, so = ba_cpu.solve(p, max_iters=iters, loss=loss)
    print(f"  engine: {sm}")
    print(f"  oracle: status={so.status} it={so.iterations} acc={so.accepted} rej={so.rejected} "
          f"c0={so.initial_cost:.9e} c1={so.final_cost:.9e} mu={so.final_mu:.6e}")
    hc = ba.dump("HIST_COST", (1, 64))[0, :iters]
    hm = ba.dump("HIST_MU", (1, 64))[0, :iters]
    ha = ba.dump("HIST_ACCEPT", (1, 64), np.int32)[0, :iters]
    print("  cost hist rel", rel(hc, np.array(so.cost_hist[:iters])), "mu hist rel",
          rel(hm, np.array(so.mu_hist[:iters])), "accept equal", bool((ha == np.array(so.accept_hist[:iters])).all()))
    print("  final poses abs", float(np.abs(poses - po).max()), "points abs", float(np.abs(pts - xo).max()),
          "rmse engine", synth.rmse_px(p, poses, pts), "oracle", synth.rmse_px(p, po, xo), f"wall {t1 - t0:.4f}s")
    ba.close()


def time_batch(nwin, iters=10, distinct=64):
    b = synth.make_batch(nwin, distinct=min(distinct, nwin))
    ba = BundleAdjuster(K_of(b), BAConfig(max_iters=iters))
    t0 = time.time()
    ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, b.cam_off, b.pt_off, b.obs_off)
    torch.cuda.synchronize()
    t_setup = time.time() - t0
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for rep in range(4):
        ba.reset()
        torch.cuda.synchronize()
        ev[0].record()
        ba.solve(wait=False)
        ev[1].record()
        torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    sms = ba.summaries()
    its = sum(s.iterations for s in sms)
    ms = min(ts[1:])
    print(f"=== batch {nwin} windows, {b.no} obs: setup {t_setup:.3f}s solve ms {ts} -> "
          f"{b.no * iters / ms * 1e3:.4e} obs*iter/s ; iterations total {its}; "
          f"status counts {np.bincount([s.status for s in sms])}", flush=True)
    if nwin <= 64:
        po, xo, sos = ba_cpu.solve_batch(b, max_iters=iters, threads=0)
        poses, pts = ba.get()
        fc = np.array([s.final_cost for s in sms])
        fo = np.array([s.final_cost for s in sos])
        print("  batch final-cost rel", rel(fc, fo), "poses abs", float(np.abs(poses - po).max()),
              "points abs", float(np.abs(pts - xo).max()))
    ba.close()
    return ms


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), "host cores", os.cpu_count())
    check_problem("tiny", synth.make_window(3, n_cams=4, n_pts=50, n_obs=150))
    check_problem("C1", synth.make_window(0))
    check_problem("C3-malaga", synth.make_window(1, n_cams=20, cam=synth.MALAGA, outlier_frac=0.1, n_obs=30000),
                  loss="huber")
    for nw in (16, 64, 592, 4096):
        time_batch(nw)
