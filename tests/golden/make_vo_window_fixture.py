"""Windows the VO loop REALLY hands to the bundle adjuster: the pinned frame loop
(visualodometry_b200/vo_loop.py == src/modules/vo.py of the reference, tests/test_vo_loop.py) runs on the synthetic
sequence with the CPU oracle as its BA, and the arrays of selected calls -- exactly what adapter.pack_window produced:
Huber a = 2, two gauge-fixed keyframes, ragged tracks, 2-observation landmarks -- are stored with the oracle's result.

    python tests/golden/make_vo_window_fixture.py        (CPU only; writes tests/golden/vo_windows.npz)
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_ba import OracleBA  # noqa: E402
from visualodometry_b200 import BAConfig, adapter, vo_loop, vo_synth  # noqa: E402
from visualodometry_b200.ba import poses_from_T  # noqa: E402

KEEP = (0, 1, 3, 8, 15, 24)      # call numbers kept: the window grows from 2 to 10 keyframes over the first calls
FRAMES = 140


def record(n_frames=FRAMES, keep=KEEP, solver_factory=None):
    seq = vo_synth.SynthSequence(n_frames=n_frames, seed=0)
    cfg = BAConfig(window=10, max_iters=10, loss="huber", loss_scale=2.0)
    cv2.setRNGSeed(0)
    vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
    solver = (solver_factory or OracleBA)(seq.K, cfg)
    inner, calls = solver.optimize, []

    def spy(T_cw, points, obs_cam, obs_pt, obs_uv, fixed_cam):
        res = inner(T_cw, points, obs_cam, obs_pt, obs_uv, fixed_cam)
        k = len(calls)
        calls.append(None)
        if keep is None or k in keep:
            calls[k] = dict(T_cw=np.array(T_cw), points=np.array(points), obs_cam=np.array(obs_cam, np.int32),
                            obs_pt=np.array(obs_pt, np.int32), obs_uv=np.array(obs_uv, np.float32),
                            fixed=np.array(fixed_cam, np.uint8), res_T=np.array(res.T_cw), res_X=np.array(res.points),
                            summary=np.array([res.summary.status, res.summary.iterations, res.summary.accepted,
                                              res.summary.rejected, res.summary.n_invalid], np.int64),
                            costs=np.array([res.summary.initial_cost, res.summary.final_cost]))
        return res
    solver.optimize = spy
    adapter.install(vo, cfg, ba=solver)
    for i in range(seq.n_frames):
        vo.process_frame(seq.image(i))
    return seq, cfg, calls


def main():
    seq, cfg, calls = record()
    out = {"K": seq.K, "n_calls": np.array(len(calls)), "keep": np.array([k for k in KEEP if k < len(calls)])}
    for k in out["keep"]:
        c = calls[int(k)]
        for name, v in c.items():
            out[f"w{k}_{name}"] = v
        print(f"call {k}: nc={c['T_cw'].shape[0]} np={c['points'].shape[0]} no={c['obs_cam'].size} fixed={int(c['fixed'].sum())} "
              f"summary={c['summary'].tolist()} cost {c['costs'][0]:.6g} -> {c['costs'][1]:.6g}")
    path = os.path.join(ROOT, "tests", "golden", "vo_windows.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", len(calls), "BA calls in", FRAMES, "frames")


if __name__ == "__main__":
    main()
