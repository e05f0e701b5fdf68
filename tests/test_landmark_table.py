"""Dense id -> slot landmark table: identical to the dict scans of the reference's frame loop
(`vo.py:121-130`, `vo.py:35-47`) -- on random operation sequences and inside the pinned loop (same poses, keyframes,
ids and map as the unmodified reference, bit for bit; tests/golden/reference_vo_traj.npz)."""
import os

import numpy as np
import pytest

from visualodometry_b200.landmark_table import LandmarkTable

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_table_equals_dict_on_random_operation_sequences():
    rng = np.random.default_rng(0)
    t, d = LandmarkTable(capacity=64), {}
    nxt = 0
    for step in range(400):
        op = rng.integers(0, 6)
        if op <= 1:                               # consecutive inserts (vo.py:281-283)
            for _ in range(int(rng.integers(1, 60))):
                x = rng.normal(size=3)
                t[nxt] = x; d[nxt] = x
                nxt += 1
        elif op == 2 and d:                       # prune below a floor (vo.py:35-47)
            floor = int(rng.integers(min(d), nxt + 1))
            t.prune_below(floor)
            for k in [k for k in d if k < floor]:
                del d[k]
        elif op == 3 and d:                       # culling pops
            for k in rng.choice(list(d), size=min(5, len(d)), replace=False).tolist():
                assert np.array_equal(t.pop(k), d.pop(k))
            assert t.pop(-7, None) is None
        elif op == 4 and d:                       # write-back of refined landmarks, some ids no longer present
            ids = rng.integers(max(0, min(d) - 5), nxt + 5, size=40)
            rows = rng.normal(size=(40, 3))
            t.update_rows(ids, rows)
            d.update((int(p), x) for p, x in zip(ids.tolist(), np.array(rows)) if p in d)
        else:                                     # the matched-id look-up of the tracker (vo.py:121-130)
            ids = rng.integers(-1, nxt + 10, size=200)
            valid, X = t.lookup(ids)
            v0 = np.array([(p != -1 and p in d) for p in ids])
            assert np.array_equal(valid, v0)
            X0 = np.array([d[p] for p in ids[v0]]).reshape(-1, 3)
            assert np.array_equal(X, X0)
        assert len(t) == len(d) and set(t) == set(d)
    for k in d:
        assert np.array_equal(t[k], d[k])
    t.clear()
    assert len(t) == 0 and not t.lookup(np.arange(nxt))[0].any()


def test_table_rebases_when_an_older_id_arrives():
    t = LandmarkTable(capacity=8)
    t[100] = np.ones(3)
    t[3] = np.zeros(3)
    t[150] = 2 * np.ones(3)
    v, X = t.lookup([3, 4, 100, 150, -1])
    assert v.tolist() == [True, False, True, True, False] and X[:, 0].tolist() == [0.0, 1.0, 2.0]


def test_pinned_loop_with_the_dense_table_is_still_the_reference_bit_for_bit():
    cv2 = pytest.importorskip("cv2")
    from visualodometry_b200 import vo_loop, vo_synth
    gold = np.load(os.path.join(ROOT, "tests", "golden", "reference_vo_traj.npz"))
    seq = vo_synth.SynthSequence(n_frames=int(gold["n_frames"]), seed=int(gold["seed"]))
    cv2.setRNGSeed(0)
    vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
    vo.dense_map = True
    poses = []
    for i in range(seq.n_frames):
        vo.process_frame(seq.image(i))
        poses.append(vo.T_wc.copy())
    assert isinstance(vo.map_points, LandmarkTable)
    tol = 0.0 if str(gold["cv2_version"]) == cv2.__version__ else 1e-9
    assert np.abs(np.array(poses) - gold["T_wc"]).max() <= tol
    assert vo.next_pt_id == int(gold["next_pt_id"])
    assert np.array_equal(np.array(sorted(vo.map_points)), gold["map_ids"])
    assert np.abs(np.array([vo.map_points[p] for p in gold["map_ids"]]) - gold["map_xyz"]).max() <= tol
    assert np.array_equal(vo.keyframe["ids"], gold["last_ids"])


def test_adapter_with_the_dense_table_equals_the_dict_path(oracle):
    """pack_window / write_back through LandmarkTable give the same arrays and the same map as through the dict."""
    cv2 = pytest.importorskip("cv2")
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_ba import OracleBA
    from visualodometry_b200 import BAConfig, adapter, vo_loop, vo_synth
    seq = vo_synth.SynthSequence(n_frames=60, seed=1)
    cfg = BAConfig(window=6, max_iters=4, loss="huber", loss_scale=2.0)
    out = []
    for dense in (False, True):
        cv2.setRNGSeed(0)
        vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
        vo.dense_map = dense
        ad = adapter.install(vo, cfg, ba=OracleBA(seq.K, cfg))
        poses = []
        for i in range(seq.n_frames):
            vo.process_frame(seq.image(i))
            poses.append(vo.T_wc.copy())
        out.append((np.array(poses), dict(vo.map_points), len(ad.summaries)))
    assert out[0][2] == out[1][2] and out[0][2] > 3
    assert np.array_equal(out[0][0], out[1][0])
    assert sorted(out[0][1]) == sorted(out[1][1])
    for k in out[0][1]:
        assert np.array_equal(out[0][1][k], out[1][1][k])
