"""The two oracles against each other (CPU): oracle/ba_numpy.py (numpy / scipy.sparse, whole-matrix algebra) and
oracle/ba_cpu.cpp (C++, block by block) are independent restatements of Spec a0 (SURVEY.md §8 row c).  They must
agree stage by stage to rounding and take the same LM trajectory.  The reference ships no BA, so this pair -- not a
reference output -- is what the LM / Schur stages of the engine are held against (parity unpinned)."""
import numpy as np
import pytest

from oracle import ba_numpy
from visualodometry_b200 import synth


def _two_fixed():
    p = synth.make_window(9, n_cams=8, n_pts=500, n_obs=3000)
    p.fixed_cam[3] = 1
    return p


PROBLEMS = {
    "tiny": (lambda: synth.make_window(3, n_cams=4, n_pts=50, n_obs=150), "trivial"),
    "two_fixed": (_two_fixed, "trivial"),
    "huber": (lambda: synth.make_window(1, n_cams=12, n_pts=400, n_obs=3000, cam=synth.MALAGA, outlier_frac=0.1), "huber"),
    "c1": (lambda: synth.make_window(0), "trivial"),
}


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


@pytest.mark.parametrize("name", list(PROBLEMS))
def test_stage_by_stage(oracle, name):
    p, loss = PROBLEMS[name][0](), PROBLEMS[name][1]
    a = ba_numpy.stage(p, mu=1e4, loss=loss)
    b = oracle.stage(p, mu=1e4, loss=loss)
    for k, tol in (("r", 1e-12), ("w", 1e-12), ("Jc", 1e-12), ("Jp", 1e-12), ("B", 1e-12), ("gc", 1e-11), ("C", 1e-12),
                   ("gp", 1e-11), ("S", 1e-10), ("gred", 1e-9), ("dc", 1e-8), ("dp", 1e-8)):
        assert rel(a[k], np.asarray(b[k]).reshape(np.shape(a[k]))) <= tol, (k, rel(a[k], np.asarray(b[k]).reshape(np.shape(a[k]))))
    assert abs(a["cost"] - b["cost"]) <= 1e-13 * b["cost"]
    assert abs(a["model_decrease"] - b["model_decrease"]) <= 1e-9 * abs(b["model_decrease"])
    assert abs(a["trial_cost"] - b["trial_cost"]) <= 1e-9 * b["trial_cost"]


@pytest.mark.parametrize("name", list(PROBLEMS))
def test_lm_trajectory(oracle, name):
    p, loss = PROBLEMS[name][0](), PROBLEMS[name][1]
    pa, xa, sa = ba_numpy.solve(p, max_iters=8, loss=loss)
    pb, xb, sb = oracle.solve(p, max_iters=8, loss=loss)
    n = sb.iterations
    assert sa.iterations == n
    # accept sequence / radius while a step still changes the cost by more than rounding
    live = [i for i in range(n) if i == 0 or abs(sb.cost_hist[i] - sb.cost_hist[i - 1]) > 1e-7 * sb.cost_hist[i]]
    for i in live:
        assert sa.accept_hist[i] == sb.accept_hist[i], i
        assert abs(sa.mu_hist[i] - sb.mu_hist[i]) <= 1e-6 * sb.mu_hist[i], i
    assert rel(sa.cost_hist[:n], list(sb.cost_hist)[:n]) <= 1e-9
    assert abs(synth.rmse_px(p, pa, xa) - synth.rmse_px(p, pb, xb)) <= 1e-6 * synth.rmse_px(p, pb, xb)
    assert np.abs(pa - pb).max() <= 1e-5


def test_pcg_solver_agrees(oracle):
    p = synth.make_window(5, n_cams=8, n_pts=300, n_obs=2000)
    a = ba_numpy.stage(p, solver="pcg")
    b = oracle.stage(p, solver="pcg")
    assert rel(a["dc"], b["dc"]) <= 1e-8
    pa, xa, sa = ba_numpy.solve(p, max_iters=5, solver="pcg")
    pb, xb, sb = oracle.solve(p, max_iters=5, solver="pcg")
    assert sa.pcg_iters_total == sb.pcg_iters_total
    assert np.abs(pa - pb).max() <= 1e-7


def _near_gate_problem():
    """A window whose landmark 0 sits just in front of camera 1 with a grossly wrong observation: the cheapest way
    to lower the cost is to push it BEHIND the gate (its observations then cost 0)."""
    p = synth.make_window(11, n_cams=5, n_pts=120, n_obs=500)
    sel = np.flatnonzero(p.obs_pt == 0)
    R = p.poses[1, :9].reshape(3, 3)
    cam1_centre = -R.T @ p.poses[1, 9:]
    p.points[0] = cam1_centre + R.T @ np.array([0.02, 0.01, 0.05])     # 5 cm in front of camera 1
    p.obs_uv[sel] += np.float32(400.0)                                   # grossly inconsistent pixels
    if not (p.obs_cam[sel] == 1).any():
        p.obs_cam[sel[0]] = 1
    return p


@pytest.mark.parametrize("impl", ["cpp", "numpy"])
def test_steps_that_discard_observations_are_rejected(oracle, impl):
    """an accepted step must not increase the number of observations behind the cheirality gate."""
    p = _near_gate_problem()
    solve = oracle.solve if impl == "cpp" else ba_numpy.solve
    base = ba_numpy.observe(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.intr, jac=False)
    n0 = int((~base["valid"]).sum())
    _, _, sm = solve(p, max_iters=15)
    assert sm.n_invalid <= n0, (sm.n_invalid, n0)
    assert sm.final_cost <= sm.initial_cost
