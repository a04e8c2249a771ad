"""Planner: the C oracle against fixtures recorded from the reference's own
attackers_policy_1 (oracle/make_golden_planner.py), and the keyed k-means against scipy's kmeans2 statistically."""
import glob
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), 'oracle'))
import dca_oracle as orc  # noqa: E402

PLANNER_CASES = sorted(os.path.basename(p)[len('planner_'):-4] for p in glob.glob(os.path.join(HERE, 'golden', 'planner_*.npz')))


def load_planner_case(name):
    return np.load(os.path.join(HERE, 'golden', 'planner_%s.npz' % name))


@pytest.mark.parametrize('name', PLANNER_CASES)
def test_oracle_planner_core_matches_reference(name):
    """DBSCAN targets, both assignments and the ring slots reproduce the reference's attacker forces when its own
    k-means result is injected: <= 1e-9 absolute (cos/sin/sqrt are glibc on both sides), on every recorded step."""
    d = load_planner_case(name)
    n_checked = 0
    for t in range(len(d['label'])):
        want = d['force'][t]
        if np.isnan(want).any():
            continue  # done step: the env was reset before the forces could be read back
        got = orc.plan_core(d['gpos'][t], d['galive'][t], d['apos'][t], d['aalive'][t], d['label'][t], d['cent'][t],
                            bool(d['disable_red']))
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-9, err_msg='%s step %d' % (name, t))
        n_checked += 1
    assert n_checked >= 40


def test_keyed_kmeans_is_a_lloyd_fixed_point_and_statistically_scipy():
    """The keyed k-means++ / Lloyd restatement: (1) structural properties of kmeans2's result, (2) its within-cluster
    SSE is distributed like scipy.cluster.vq.kmeans2's on the same data (mean within 3 %)."""
    from scipy.cluster.vq import kmeans2
    rng = np.random.RandomState(5)
    sse_ours, sse_ref = [], []
    for trial in range(60):
        n, k = 50, 10
        pts = np.stack([2 + 0.3 * rng.randn(n), np.linspace(-4, 4, n) * rng.choice([0.5, 0.75, 1.0], n)], 1) + 0.05 * rng.randn(n, 2)
        label, cent, raised = orc.kmeans(pts, k, seed=77, gid=trial, counter=3)
        assert label.min() == 0 and label.max() == k - 1 and len(np.unique(label)) == k
        for c in range(k):  # centroids are the sequential means of the returned labels
            m = pts[label == c]
            acc = np.zeros(2)
            for row in m:
                acc = acc + row
            assert np.array_equal(cent[c], acc / len(m))
        sse_ours.append(((pts - cent[label]) ** 2).sum())
        for s in range(100):
            try:
                c2, l2 = kmeans2(pts, k, iter=20, minit='++', seed=s + 1000 * trial, missing='raise')
                break
            except Exception:
                continue
        sse_ref.append(((pts - c2[l2]) ** 2).sum())
    a, b = np.mean(sse_ours), np.mean(sse_ref)
    assert abs(a - b) / b < 0.03, (a, b)
    # deterministic in its key
    l1, c1, _ = orc.kmeans(pts, 10, 77, 5, 9)
    l2, c2, _ = orc.kmeans(pts, 10, 77, 5, 9)
    assert np.array_equal(l1, l2) and np.array_equal(c1, c2)
