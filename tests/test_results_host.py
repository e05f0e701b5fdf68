"""Host-side result objects of the Python mirror (visualodometry_b200/ba.py): BAResult builds its 4 x 4 matrices on first
access, SummaryList is a sequence view of the C array a batch call filled.  CPU only (no engine call)."""
import numpy as np

from visualodometry_b200 import _lib
from visualodometry_b200.ba import BAResult, BASummary, SummaryList, T_from_poses, poses_from_T


def test_baresult_builds_matrices_lazily_and_round_trips():
    rng = np.random.default_rng(0)
    poses = rng.normal(size=(5, 12))
    r = BAResult(None, np.zeros((3, 3)), None, poses=poses)
    assert r._T is None
    T = r.T_cw
    assert T.shape == (5, 4, 4) and np.array_equal(T, T_from_poses(poses)) and r.T_cw is T
    assert np.array_equal(poses_from_T(T), poses)
    # the positional form the adapter tests and the oracle wrapper use
    sm = BASummary(1, 10, 9, 1, 0, 0, 2.0, 1.0, 1e4)
    r2 = BAResult(T, np.ones((3, 3)), sm, [sm])
    assert r2.T_cw is T and r2.summary is sm and r2.summaries == [sm] and np.array_equal(r2.poses, poses)


def test_summary_list_is_a_view_of_the_c_array():
    n = 7
    arr = (_lib.Summary * n)()
    for i in range(n):
        arr[i].status, arr[i].iterations, arr[i].accepted, arr[i].final_cost = (3 if i == 4 else 1), 10, i, float(i)
    s = SummaryList(arr)
    assert len(s) == n and s.array.dtype.itemsize == 48 and s.array["accepted"].tolist() == list(range(n))
    assert isinstance(s[2], BASummary) and s[2].accepted == 2 and s[-1].final_cost == 6.0
    assert [x.accepted for x in s] == list(range(n)) and [x.accepted for x in s[1:3]] == [1, 2]
    w = s.worst()
    assert w.status == 3 and w.accepted == 4 and not w.ok and s[0].ok
    assert len(SummaryList((_lib.Summary * 0)())) == 0
