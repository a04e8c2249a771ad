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


# ------------------------------------------------------------------------------------------------ GPU (run with -m gpu)
def _blank_state(E, G, A, gpos, galive, apos, aalive):
    N = G + A
    st = dict(pos=np.concatenate([gpos, apos], 1), vel=np.zeros((E, N, 2)), atk_rad=np.zeros((E, N)),
              terrain=np.ones((E, N)), alive=np.concatenate([galive, aalive], 1).astype(np.uint8),
              just_died=np.zeros((E, N), np.uint8), hit=np.zeros((E, N), np.uint8), was_hit=np.zeros((E, N), np.uint8),
              num_hit=np.zeros((E, N), np.int32), num_was_hit=np.zeros((E, N), np.int32), init_theta=np.zeros((E,)),
              time_step=np.zeros((E,), np.int32), n_alive_guards=galive.sum(1).astype(np.int32),
              n_alive_attackers=aalive.sum(1).astype(np.int32), obs_counter=np.zeros((E,), np.uint32),
              episode=np.zeros((E,), np.uint32))
    return st


@pytest.mark.gpu
@pytest.mark.parametrize('name', PLANNER_CASES)
def test_cuda_planner_core_matches_reference_and_oracle(name):
    """k_plan with the reference's own k-means result injected: every recorded step of the fixture as one env of a
    batch.  Forces vs the reference <= 1e-9 absolute, vs the C oracle <= 1e-12 (CUDA vs glibc cos/sin/sqrt)."""
    import torch
    from test_gpu_parity import make_gpu_env
    d = load_planner_case(name)
    G, A, E = int(d['G']), int(d['A']), len(d['label'])
    cfg = dict(num_guards=G, num_attackers=A, DISALBE_RED_FUNCTION=bool(d['disable_red']))
    env = make_gpu_env(cfg, E, seed=1, gid0=0, debug=False)
    env.reset()
    env.import_state(_blank_state(E, G, A, d['gpos'], d['galive'], d['apos'], d['aalive']))
    lab = torch.from_numpy(d['label'].astype(np.int32)).to(env.device)
    cen = torch.from_numpy(np.ascontiguousarray(d['cent'], np.float64)).to(env.device)
    got = env.plan(team_label=lab, team_centroid=cen).cpu().numpy()
    n_checked = 0
    for t in range(E):
        want_orc = orc.plan_core(d['gpos'][t], d['galive'][t], d['apos'][t], d['aalive'][t], d['label'][t], d['cent'][t],
                                 bool(d['disable_red']))
        np.testing.assert_allclose(got[t], want_orc, rtol=0, atol=1e-12, err_msg='%s step %d vs oracle' % (name, t))
        if not np.isnan(d['force'][t]).any():
            np.testing.assert_allclose(got[t], d['force'][t], rtol=0, atol=1e-9, err_msg='%s step %d vs reference' % (name, t))
            n_checked += 1
    assert n_checked >= 40


@pytest.mark.gpu
@pytest.mark.parametrize('cfg,E,steps', [
    (dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True, terrain_parameters=[0.05, 0.2],
          MaxEpisodeStep=60, init_distance=2), 96, 80),
    (dict(num_guards=20, num_attackers=33, introduce_terrain=False, MaxEpisodeStep=25, init_distance=1.5), 40, 60),
])
def test_cuda_planner_native_matches_oracle_in_a_rollout(cfg, E, steps):
    """Full planner (keyed k-means++ teams with their persistent state) GPU vs C oracle on identical states over a
    rollout with deaths and autoresets; the step itself is driven by the oracle's forces and re-synced, so any
    difference is the planner's.  Forces <= 1e-12 absolute (a different team label would show as O(1))."""
    import torch
    from dca_oracle import OraclePlanner, OracleVecEnv
    from test_gpu_parity import make_gpu_env
    o = OracleVecEnv(cfg, E, seed=5, env_gid0=100, nthreads=8)
    env = make_gpu_env(cfg, E, seed=5, gid0=100, debug=False)
    o.reset(); env.reset()
    env.import_state(o.state)
    planner = OraclePlanner(o)
    rng = np.random.RandomState(3)
    n_done = 0
    for t in range(steps):
        want = planner.plan()
        got = env.plan().cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-12, err_msg='t=%d' % t)
        act = rng.randint(0, 7, size=(E, o.G))
        o.step(act.astype(np.float64), want)
        n_done += int((o.done != 0).sum())
        env.import_state(o.state)
    assert n_done > 0
    assert np.abs(want).max() > 0.1


@pytest.mark.gpu
def test_keyed_planner_episode_statistics_match_reference():
    """The one statistical sub-step: the k-means++ team draws come from the keyed stream, not from
    numpy's MT19937, so team composition is the reference's in distribution only.  640 episodes of the unmodified
    reference env with its own sklearn / scipy planner under uniform random guard actions
    (tests/golden/planner_stats_50v50.json, oracle/make_golden_planner_stats.py) against 640 episodes of the GPU env with
    k_plan (320 envs x their first two episodes): episode length, guard win rate, share of time-limit endings and
    guards alive at the end within sampling error."""
    import json
    import torch
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.vec_env import DcaVecEnv
    ref = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'planner_stats_50v50.json')))
    assert ref['episodes'] >= 500
    scfg = type('S', (ScenarioConfig,), dict(num_guards=50, num_attackers=50, size=5.0, random_jam_prob=0.05,
                                             introduce_terrain=True, terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180,
                                             half_death_reward=True, init_distance=4))
    E, per_env = 320, 2
    env = DcaVecEnv(E, scfg=scfg, seed=31337)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(5)
    ep = np.zeros(E, int); cur = np.zeros(E, int)
    lens, wins, alive_end = [], [], []
    for t in range(per_env * 181 + 2):
        act = torch.randint(0, 7, (E, 50), generator=g, device=env.device, dtype=torch.int32)
        env.step(act)   # plan() inside: the scripted attackers
        done = env.done.cpu().numpy().astype(bool)
        cur += 1
        if done.any():
            win = env.win.cpu().numpy()
            tob = env.tobs.cpu().numpy()
            for e in np.nonzero(done)[0]:
                if ep[e] < per_env:
                    lens.append(cur[e]); wins.append(bool(win[e])); alive_end.append(int((tob[e, :, 0, 0] > 0).sum()))
                ep[e] += 1; cur[e] = 0
        if (ep >= per_env).all():
            break
    n, n_ref = len(lens), ref['episodes']
    assert n == E * per_env
    mean_len, tl = float(np.mean(lens)), float(np.mean(np.array(lens) >= 180))
    print('keyed planner: %d episodes, length %.1f +- %.1f (reference %.1f +- %.1f), time-limit endings %.3f (%.3f), guard '
          'wins %.3f (%.3f), guards alive at the end %.3f (%.3f)' % (
              n, mean_len, np.std(lens), ref['mean_len'], ref['std_len'], tl, ref['frac_time_limit'], np.mean(wins),
              ref['guard_win_rate'], np.mean(alive_end), ref['mean_guards_alive_at_end']))
    se_len = np.sqrt(np.var(lens) / n + ref['std_len'] ** 2 / n_ref)
    assert abs(mean_len - ref['mean_len']) <= 4 * se_len + 1.0
    p = 0.5 * (tl + ref['frac_time_limit'])
    assert abs(tl - ref['frac_time_limit']) <= 4 * np.sqrt(max(p * (1 - p), 1e-4) * (1.0 / n + 1.0 / n_ref)) + 0.01
    assert abs(np.mean(wins) - ref['guard_win_rate']) <= 0.02
    se_a = np.sqrt(np.var(alive_end) / n + ref['std_guards_alive_at_end'] ** 2 / n_ref)
    assert abs(np.mean(alive_end) - ref['mean_guards_alive_at_end']) <= 4 * se_a + 0.05


@pytest.mark.gpu
@pytest.mark.parametrize('G,A', [(50, 50), (40, 90), (100, 100)])
def test_cuda_planner_every_team_size_path_matches_oracle(G, A):
    """Injected team labels that force every member->slot path of k_plan on random states -- teams of 1..8 (subset DP),
    9..16 (register assignment, precomputed costs), 17..32 (register assignment), more than 32 members (shared-memory
    assignment), empty teams -- and, at 100v100, the workspaces that share one region of shared memory at their largest:
    forces vs the C oracle (augmenting paths throughout) <= 1e-12."""
    import torch
    from test_gpu_parity import make_gpu_env
    rng = np.random.RandomState(G * 1000 + A)
    K = A // 5
    E = 48
    gpos = rng.uniform(-4, 4, size=(E, G, 2)); apos = rng.uniform(-4, 4, size=(E, A, 2))
    galive = rng.rand(E, G) < 0.8; galive[:, 0] = True
    aalive = rng.rand(E, A) < 0.85
    label = np.zeros((E, A), np.int32)
    for e in range(E):
        mode = e % 6
        if mode == 0:      # one team holds everybody (> 32 members when A > 32)
            label[e] = 0
        elif mode == 1:    # a team of 17..32, the rest spread
            n_big = rng.randint(17, min(32, A - 1) + 1)
            label[e, :n_big] = 1
            label[e, n_big:] = rng.randint(2, K, size=A - n_big)
        elif mode == 2:    # teams of 9..16
            sizes = []
            while sum(sizes) < A:
                sizes.append(min(rng.randint(9, 17), A - sum(sizes)))
            label[e] = np.repeat(np.arange(len(sizes)) % K, sizes)[:A]
        elif mode == 3:    # small teams only (round robin): sizes about 5
            label[e] = rng.permutation(A) % K
        elif mode == 4:    # random sizes incl. singletons and empty teams
            label[e] = rng.randint(0, max(2, K // 2), size=A)
        else:              # one team of more than 32 and small ones
            n_big = min(A - 3, 33 + rng.randint(0, 8))
            label[e, :n_big] = K - 1
            label[e, n_big:] = rng.randint(0, K - 1, size=A - n_big)
        label[e] = label[e][rng.permutation(A)]
    cent = np.zeros((E, K, 2))
    for e in range(E):
        for k in range(K):
            m = label[e] == k
            cent[e, k] = apos[e, m].mean(0) if m.any() else rng.uniform(-4, 4, size=2)
    cfg = dict(num_guards=G, num_attackers=A)
    env = make_gpu_env(cfg, E, seed=1, gid0=0, debug=False)
    env.reset()
    env.import_state(_blank_state(E, G, A, gpos, galive, apos, aalive))
    got = env.plan(team_label=torch.from_numpy(label).to(env.device),
                   team_centroid=torch.from_numpy(cent).to(env.device)).cpu().numpy()
    for e in range(E):
        want = orc.plan_core(gpos[e], galive[e], apos[e], aalive[e], label[e], cent[e], False)
        np.testing.assert_allclose(got[e], want, rtol=0, atol=1e-12, err_msg='G=%d A=%d env %d (mode %d)' % (G, A, e, e % 6))
    assert np.abs(got).max() > 0.1
