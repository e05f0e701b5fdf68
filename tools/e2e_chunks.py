"""C4 end-to-end (b2ba_optimize_batch, pinned host buffers) for several chunk plans: B2BA_CHUNK_PLAN=first,step in
quarter waves.  usage: python tools/e2e_chunks.py [windows]"""
import os
import subprocess
import sys

CODE = r'''
import sys, time, numpy as np, torch
sys.path.insert(0, ".")
import bench
from types import SimpleNamespace
from visualodometry_b200 import BAConfig, BundleAdjuster, synth
nw = int(sys.argv[1])
b = synth.make_batch(nw, seed0=0, distinct=64)
ba = BundleAdjuster(bench.K_of(b.intr), BAConfig(max_iters=10))
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
hb = SimpleNamespace(poses=pin(b.poses), points=pin(b.points), obs_cam=pin(b.obs_cam), obs_pt=pin(b.obs_pt),
                     obs_uv=pin(b.obs_uv), fixed_cam=pin(b.fixed_cam), cam_off=b.cam_off, pt_off=b.pt_off, obs_off=b.obs_off)
out = (pin(np.empty_like(b.poses, dtype=np.float64)), pin(np.empty_like(b.points, dtype=np.float64)))
ba.optimize_batch(hb, out=out)
torch.cuda.synchronize()
ts = []
for _ in range(4):
    t = time.perf_counter(); ba.optimize_batch(hb, out=out); torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
print("%.2f ms (min %.2f)" % (1e3 * np.median(ts), 1e3 * min(ts)))
'''
nw = sys.argv[1] if len(sys.argv) > 1 else "4096"
for plan in ("4,8", "2,8", "1,8", "2,4", "4,4", "1,4", "2,6", "4,12", "2,2"):
    env = dict(os.environ, B2BA_CHUNK_PLAN=plan)
    r = subprocess.run([sys.executable, "-c", CODE, nw], env=env, capture_output=True, text=True)
    print(plan, r.stdout.strip() or r.stderr.strip()[-300:], flush=True)
