import sys, time, numpy as np, torch
sys.path.insert(0, ".")
import bench
from types import SimpleNamespace
from visualodometry_b200 import BAConfig, BundleAdjuster, synth
b = synth.make_batch(4096, seed0=0, distinct=64)
ba = BundleAdjuster(bench.K_of(b.intr), BAConfig(max_iters=10))
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
hb = SimpleNamespace(poses=pin(b.poses), points=pin(b.points), obs_cam=pin(b.obs_cam), obs_pt=pin(b.obs_pt),
                     obs_uv=pin(b.obs_uv), fixed_cam=pin(b.fixed_cam), cam_off=b.cam_off, pt_off=b.pt_off, obs_off=b.obs_off)
out = (pin(np.empty_like(b.poses, dtype=np.float64)), pin(np.empty_like(b.points, dtype=np.float64)))
for _ in range(3):
    t = time.perf_counter(); ba.optimize_batch(hb, out=out); torch.cuda.synchronize(); print("call %.2f ms" % (1e3*(time.perf_counter()-t)), file=sys.stderr)
