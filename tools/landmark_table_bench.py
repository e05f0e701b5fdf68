"""Host-side cost of the reference's dict scans against the dense id -> slot table, CPU only:
the pinned frame loop (no BA) with `dense_map` off / on, and the two look-up forms on the final map in isolation.
    python tools/landmark_table_bench.py [frames]  -> one JSON line"""
import json
import os
import sys
import time

import cv2
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from visualodometry_b200 import vo_loop, vo_synth  # noqa: E402
from visualodometry_b200.landmark_table import LandmarkTable  # noqa: E402


def run(seq, dense):
    cv2.setRNGSeed(0)
    vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
    vo.dense_map = dense
    t = time.perf_counter()
    for i in range(seq.n_frames):
        vo.process_frame(seq.image(i))
    return time.perf_counter() - t, vo


def main():
    frames = int(sys.argv[1]) if len(sys.argv) > 1 else 600
    seq = vo_synth.SynthSequence(n_frames=frames, seed=0)
    t_dict, vo = run(seq, False)
    t_dense, vo2 = run(seq, True)
    same = bool(np.array_equal(np.array(vo.trajectory), np.array(vo2.trajectory)))
    # the look-up of vo.py:121-130 in isolation, on the final map and the last keyframe's ids
    mp, ids = vo.map_points, vo.keyframe["ids"]
    table = LandmarkTable(mp)
    reps = 200
    t = time.perf_counter()
    for _ in range(reps):
        valid = np.array([(p != -1 and p in mp) for p in ids])
        X = np.array([mp[p] for p in ids[valid]])
    t_scan = (time.perf_counter() - t) / reps
    t = time.perf_counter()
    for _ in range(reps):
        v2, X2 = table.lookup(ids)
    t_tab = (time.perf_counter() - t) / reps
    assert np.array_equal(valid, v2) and np.array_equal(X.reshape(-1, 3), X2)
    print(json.dumps({"frames": frames, "loop_s_dict": t_dict, "loop_s_dense_table": t_dense, "same_trajectory": same,
                      "lookup_us_dict_scan": 1e6 * t_scan, "lookup_us_table": 1e6 * t_tab, "ids_per_lookup": int(ids.size),
                      "map_size": len(mp), "note": "no BA in the loop; the loop is bound by cv2.solvePnPRansac"}))


if __name__ == "__main__":
    main()
