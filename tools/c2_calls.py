"""Per-call BA time inside the C2 VO loop (which calls are slow?)."""
import time, sys
import numpy as np
import cv2
from visualodometry_b200 import BAConfig, BundleAdjuster, adapter, vo_loop, vo_synth
seq = vo_synth.SynthSequence(n_frames=int(sys.argv[1]) if len(sys.argv) > 1 else 400, seed=0)
cfg = BAConfig(window=10, max_iters=10, loss="huber", loss_scale=2.0)
cv2.setRNGSeed(0)
vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
solver = BundleAdjuster(seq.K, cfg)
inner = solver.optimize
calls = []
def timed(*a):
    t = time.perf_counter(); r = inner(*a); dt = time.perf_counter() - t
    calls.append((dt, a[0].shape[0], a[1].shape[0], len(a[2]), int(np.sum(a[5])), r.summary.status, r.summary.iterations))
    return r
solver.optimize = timed
adapter.install(vo, cfg, ba=solver)
for i in range(seq.n_frames):
    vo.process_frame(seq.image(i))
c = np.array([x[0] for x in calls])
print("calls", len(c), "total", c.sum(), "median ms", 1e3 * np.median(c), "mean ms", 1e3 * c.mean())
for x in sorted(calls, reverse=True)[:8]:
    print("  %.2f ms  nc=%d np=%d no=%d fixed=%d status=%d iters=%d" % (1e3 * x[0], *x[1:]))
for x in calls[:6]:
    print("  first: %.2f ms  nc=%d np=%d no=%d fixed=%d status=%d iters=%d" % (1e3 * x[0], *x[1:]))
