"""Developer diagnostic (GPU box): where the end-to-end time of set_problem / solve / get goes."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from visualodometry_b200 import BAConfig, BundleAdjuster, synth  # noqa: E402


def K_of(intr):
    return np.array([[intr[0], 0, intr[2]], [0, intr[1], intr[3]], [0, 0, 1]])


def run(name, b, reps=4, batch=True):
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    hb = dict(poses=pin(b.poses), points=pin(b.points), obs_cam=pin(b.obs_cam), obs_pt=pin(b.obs_pt),
              obs_uv=pin(b.obs_uv), fixed=pin(b.fixed_cam))
    out = (pin(np.empty_like(b.poses)), pin(np.empty_like(b.points)))
    offs = (b.cam_off, b.pt_off, b.obs_off) if batch else (None, None, None)
    ba = BundleAdjuster(K_of(b.intr), BAConfig(max_iters=10))
    print(f"== {name}: windows {b.nb} obs {b.no}")
    for r in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ba.set_problem(hb["poses"], hb["points"], hb["obs_cam"], hb["obs_pt"], hb["obs_uv"], hb["fixed"], *offs)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        ba.solve()
        t2 = time.perf_counter()
        if r % 2 == 0:
            ba.get()
            kind = "pageable"
        else:
            ba.get(out=out)
            kind = "pinned"
        t3 = time.perf_counter()
        print(f"  rep {r}: set_problem {1e3 * (t1 - t0):8.2f} ms  solve {1e3 * (t2 - t1):8.2f} ms  "
              f"get[{kind}] {1e3 * (t3 - t2):8.2f} ms  total {1e3 * (t3 - t0):8.2f} ms")
    ba.close()


if __name__ == "__main__":
    run("C1", synth.pack_batch([synth.make_window(0)]), reps=6)
    run("C1 single-problem API", synth.pack_batch([synth.make_window(0)]), reps=4, batch=False)
    run("C3", synth.pack_batch([synth.make_window(1, n_cams=20, cam=synth.MALAGA, outlier_frac=0.1, n_obs=30000)]))
    run("C4/8", synth.make_batch(512, seed0=0, distinct=64))
    run("C4", synth.make_batch(4096, seed0=0, distinct=128))
