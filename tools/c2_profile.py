"""cProfile of the C2 loop with every GPU stage on (BA, PnP consensus, triangulation).  usage: python tools/c2_profile.py [frames]"""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402
import numpy as np  # noqa: E402

from visualodometry_b200 import BAConfig, BundleAdjuster, adapter, vo_loop, vo_synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 600
seq = vo_synth.SynthSequence(n_frames=n, seed=0)
cfg = BAConfig(window=10, max_iters=10, loss="huber", loss_scale=2.0)


def run():
    cv2.setRNGSeed(0)
    lcfg = vo_loop.LoopConfig()
    vo = vo_loop.VOLoop(seq.K, lcfg, vo_synth.SynthFrontend(seq))
    trk = BundleAdjuster(seq.K, BAConfig())
    vo.pnp_solver = lambda X, uv, T: trk.ransac_pose(T, X, uv, lcfg.pnp_reproj_err, n_hyp=1024, seed=vo.frame_id)[:3]
    vo.triangulator = lambda T1, T2, p1, p2, K, c: trk.triangulate(T1, T2, p1, p2, c.min_depth, c.max_reproj_err)
    adapter.install(vo, cfg, ba=BundleAdjuster(seq.K, cfg))
    for i in range(seq.n_frames):
        vo.process_frame(seq.image(i))


run()
pr = cProfile.Profile()
pr.enable()
run()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
