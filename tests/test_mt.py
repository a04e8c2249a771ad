"""dca_multiteam: the oracle restatement (oracle/mt_oracle.c) and the CUDA path against trajectories
recorded from the reference's own env under the keyed RNG stream (tests/golden/mt_*.npz, oracle/make_golden_mt.py).
Bars as for DCA: integer / index outputs bit-exact, floats within 1e-5 relative."""
import glob
import json
import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, 'oracle'))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import golden_util as gu  # noqa: E402

GOLDEN_DIR = os.path.join(REPO, 'tests', 'golden')
CASES = sorted(os.path.basename(p)[3:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, 'mt_*.npz'))
               if not p.endswith('_unity.npz'))  # the team-level interface has its own test
RTOL = 1e-5
INT_KEYS = ('alive', 'just_died', 'hit', 'was_hit', 'bsa', 'time_step', 'obs_counter', 'episode')
FLOAT_KEYS = ('pos', 'vel', 'atk_rad', 'terrain', 'init_theta')
STATE_KEYS = INT_KEYS + FLOAT_KEYS


def load_case(name):
    z = np.load(os.path.join(GOLDEN_DIR, 'mt_%s.npz' % name), allow_pickle=False)
    d = {k: z[k] for k in z.files}
    d['cfg'] = json.loads(str(d['cfg_json']))
    d['seed'], d['gid'] = int(d['seed']), int(d['gid'])
    return d


def slot_ids(obs):
    """[..., N, 204] -> [..., N, 12] entity uid per row, -1 for zero rows (col 0 = alive = 1 on gathered rows, col 6 = iden)"""
    o = obs.reshape(obs.shape[:-1] + (12, 17))
    return np.where(o[..., 0] != 0, o[..., 6].astype(np.int32), -1)


def compare_obs(tag, obs, ref_obs, ref_pos):
    """Row-wise: same entity -> values within tolerance; different entities -> only acceptable when both are at exactly
    the same float32 distance from the observer.  The reference orders equal distances by np.argsort's default
    (unstable, SIMD-dispatch dependent) sort; the regular start formation (envs/collective_assault_env.py:62-70, no
    jitter) makes such ties common in the first steps of an episode.  Ours orders them by index."""
    a, b = slot_ids(obs), slot_ids(ref_obs)
    N = a.shape[0]
    p = ref_pos.astype(np.float32)
    o = obs.reshape(N, 12, 17)
    r = ref_obs.reshape(N, 12, 17)
    n_tie = 0
    for g, row in zip(*np.nonzero(a != b)):
        ia, ib = int(a[g, row]), int(b[g, row])
        assert ia >= 0 and ib >= 0, (tag, g, row, ia, ib)

        def dist(j):
            dx, dy = p[j, 0] - p[g, 0], p[j, 1] - p[g, 1]
            return np.sqrt(np.float32(dx * dx) + np.float32(dy * dy), dtype=np.float32)
        assert dist(ia) == dist(ib), (tag, g, row, ia, ib, dist(ia), dist(ib))
        n_tie += 1
    same = (a == b)[..., None] & np.ones((1, 1, 17), bool)
    gu.assert_close_rel(np.where(same, o, 0), np.where(same, r, 0), RTOL, tag + ' obs', atol=1e-12)
    return n_tie


def check_step(tag, state, obs, reward, done, ranking, case, t):
    for k in INT_KEYS:
        a = np.asarray(state[k]).astype(np.int64) & 0xFFFFFFFF
        b = np.asarray(case[k][t]).astype(np.int64) & 0xFFFFFFFF
        assert np.array_equal(a, b), (tag, k)
    for k in FLOAT_KEYS:
        gu.assert_close_rel(state[k], case[k][t], RTOL, '%s %s' % (tag, k), atol=1e-12)
    assert int(done) == int(case['done'][t]), tag
    if case['done'][t]:
        assert np.array_equal(np.asarray(ranking, np.int32), case['ranking'][t]), tag
    assert np.array_equal(np.asarray(reward, np.float64), case['reward'][t]), tag
    # the observed positions: post-reset state where the env was reset on this step (obs = reset obs)
    return compare_obs(tag, obs, case['obs'][t], case['pos'][t])


def test_golden_cases_present():
    assert set(CASES) >= {'2x25', '3teams', '4x6'}


@pytest.mark.parametrize('name', CASES)
def test_mt_oracle_matches_reference_trajectory(name):
    from mt_oracle import OracleMtEnv
    case = load_case(name)
    env = OracleMtEnv(case['cfg'], 1, seed=case['seed'], env_gid0=case['gid'])
    env.import_state(0, {k: case['s0_' + k] for k in STATE_KEYS})
    want_tobs = {int(t): case['tobs'][k] for k, t in enumerate(case['tobs_steps'])}
    n_tie = n_rows = 0
    for t in range(case['actions'].shape[0]):
        obs, rew, done, ranking = env.step(case['actions'][t][None])
        st = {k: env.state[k][0] for k in STATE_KEYS}
        n_tie += check_step('%s t=%d' % (name, t), st, obs[0], rew[0], done[0], ranking[0], case, t)
        n_rows += obs[0].size // 17
        if t in want_tobs:  # terminal obs: every row that is present in both must agree
            ta, tb = slot_ids(env.tobs[0]), slot_ids(want_tobs[t])
            assert np.array_equal(ta >= 0, tb >= 0), (name, t)
            same = (ta == tb)[..., None] & np.ones((1, 1, 17), bool)
            gu.assert_close_rel(np.where(same, env.tobs[0].reshape(-1, 12, 17), 0),
                                np.where(same, want_tobs[t].reshape(-1, 12, 17), 0), RTOL, 'tobs', atol=1e-12)
    print('%s: %d of %d obs rows differ from the reference only by the order of exactly tied distances' % (name, n_tie, n_rows))
    assert n_tie < 0.02 * n_rows


@pytest.mark.parametrize('name', CASES)
def test_mt_oracle_reset_matches_reference(name):
    from mt_oracle import OracleMtEnv
    case = load_case(name)
    env = OracleMtEnv(case['cfg'], 1, seed=case['seed'], env_gid0=case['gid'])
    obs = env.reset()
    for k in INT_KEYS:
        if k == 'just_died':
            continue
        assert np.array_equal(np.asarray(env.state[k][0]).astype(np.int64) & 0xFFFFFFFF,
                              np.asarray(case['s0_' + k]).astype(np.int64) & 0xFFFFFFFF), k
    for k in FLOAT_KEYS:
        gu.assert_close_rel(env.state[k][0], case['s0_' + k], 1e-12, 'reset ' + k, atol=1e-14)
    compare_obs('reset obs', obs[0], case['obs0'], case['s0_pos'])


# ------------------------------------------------------------------------------------------------ CUDA path
def _mt_scfg(cfg):
    from hmp2g_b200.multiteam import MtScenarioConfig
    c = type('CaseMtScenarioConfig', (MtScenarioConfig,), {})
    for k, v in cfg.items():
        assert hasattr(MtScenarioConfig, k), k
        setattr(c, k, v)
    return c


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_cuda_mt_matches_reference_golden(name):
    import torch
    assert torch.cuda.is_available()
    from hmp2g_b200.multiteam import DcaMultiTeamVecEnv
    case = load_case(name)
    env = DcaMultiTeamVecEnv(2, scfg=_mt_scfg(case['cfg']), seed=case['seed'], env_gid0=case['gid'], debug_outputs=True)
    s0 = {k: case['s0_' + k] for k in STATE_KEYS}
    env.import_state({k: np.stack([v, v]) for k, v in s0.items()})  # env 1: a bystander with another global id
    want_tobs = {int(t): case['tobs'][k] for k, t in enumerate(case['tobs_steps'])}
    dev = env.device
    n_tie = n_rows = 0
    for t in range(case['actions'].shape[0]):
        a = case['actions'][t]
        skip = bool(np.isnan(a).any())
        act = torch.from_numpy(np.tile(np.where(np.isnan(a), 0, a).astype(np.int32), (2, 1))).to(dev)
        env.step(act, skip=torch.full((2,), int(skip), dtype=torch.uint8, device=dev))
        st = {k: v[0] for k, v in env.export_state().items()}
        obs = env.obs[0].reshape(env.N, 204).cpu().numpy()
        tag = '%s t=%d' % (name, t)
        n_tie += check_step(tag, st, obs, env.reward[0].cpu().numpy(), int(env.done[0]), env.ranking[0].cpu().numpy(),
                            case, t)
        n_rows += obs.size // 17
        assert np.array_equal(slot_ids(obs), env.slot_id[0].cpu().numpy().astype(np.int32)), tag
        assert np.array_equal(env.alive[0].cpu().numpy(), case['alive'][t]), tag
        if t in want_tobs:
            tobs = env.tobs[0].reshape(env.N, 204).cpu().numpy()
            ta, tb = slot_ids(tobs), slot_ids(want_tobs[t])
            assert np.array_equal(ta >= 0, tb >= 0), tag
            same = (ta == tb)[..., None] & np.ones((1, 1, 17), bool)
            gu.assert_close_rel(np.where(same, tobs.reshape(-1, 12, 17), 0),
                                np.where(same, want_tobs[t].reshape(-1, 12, 17), 0), RTOL, 'tobs', atol=1e-12)
    assert n_tie < 0.02 * n_rows


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_cuda_mt_reset_matches_reference(name):
    import torch
    from hmp2g_b200.multiteam import DcaMultiTeamVecEnv
    case = load_case(name)
    env = DcaMultiTeamVecEnv(1, scfg=_mt_scfg(case['cfg']), seed=case['seed'], env_gid0=case['gid'])
    env.reset()
    st = env.export_state()
    for k in INT_KEYS:
        if k == 'just_died':
            continue
        assert np.array_equal(np.asarray(st[k][0]).astype(np.int64) & 0xFFFFFFFF,
                              np.asarray(case['s0_' + k]).astype(np.int64) & 0xFFFFFFFF), k
    for k in FLOAT_KEYS:
        gu.assert_close_rel(st[k][0], case['s0_' + k], 1e-9, 'reset ' + k, atol=1e-12)
    compare_obs('reset obs', env.obs[0].reshape(env.N, 204).cpu().numpy(), case['obs0'], case['s0_pos'])


@pytest.mark.gpu
@pytest.mark.parametrize('cfg,E,steps', [
    (dict(N_AGENT_EACH_TEAM=[25, 25], random_jam_prob=0.05, introduce_terrain=True, terrain_parameters=[0.05, 0.2],
          MaxEpisodeStep=40, StepsToNarrowCircle=25, init_distance=1.5), 256, 100),
    (dict(N_AGENT_EACH_TEAM=[6, 30, 11, 8], random_jam_prob=0.3, MaxEpisodeStep=25, StepsToNarrowCircle=10, size=3.0,
          init_distance=1.0, half_death_reward=False), 64, 80),
    (dict(N_AGENT_EACH_TEAM=[100, 100], introduce_terrain=True, terrain_parameters=[0.15, 0.2], MaxEpisodeStep=30,
          StepsToNarrowCircle=20, init_distance=2.0), 16, 45),
])
def test_cuda_mt_vs_oracle_fresh_rollouts(cfg, E, steps):
    """GPU and oracle side by side from their own resets, no resync: integer outputs bit-exact (alive, flags, done,
    ranking, rewards), floats within 1e-5, observation rows equal up to exactly tied distances."""
    import torch
    from hmp2g_b200.multiteam import DcaMultiTeamVecEnv
    from mt_oracle import OracleMtEnv
    seed, gid0 = 77, 1000
    gpu = DcaMultiTeamVecEnv(E, scfg=_mt_scfg(cfg), seed=seed, env_gid0=gid0, debug_outputs=True)
    orc = OracleMtEnv(cfg, E, seed=seed, env_gid0=gid0)
    gpu.reset()
    orc.reset()
    N = gpu.N
    rng = np.random.RandomState(5)
    n_done = n_tie = 0
    for t in range(steps):
        a = rng.randint(0, 7, size=(E, N))
        gpu.step(torch.from_numpy(a.astype(np.int32)).to(gpu.device))
        obs_o, rew_o, done_o, rank_o = orc.step(a.astype(np.float64))
        st = gpu.export_state()
        for k in INT_KEYS:
            assert np.array_equal(st[k].astype(np.int64) & 0xFFFFFFFF, orc.state[k].astype(np.int64) & 0xFFFFFFFF), (t, k)
        for k in FLOAT_KEYS:
            gu.assert_close_rel(st[k], orc.state[k], RTOL, 't=%d %s' % (t, k), atol=1e-12)
        done_g = gpu.done.cpu().numpy()
        assert np.array_equal(done_g, done_o), t
        d = done_o.astype(bool)
        assert np.array_equal(gpu.ranking.cpu().numpy()[d], rank_o[d]), t
        assert np.array_equal(gpu.reward.cpu().numpy().astype(np.float64), rew_o), t
        n_done += int(d.sum())
        obs_g = gpu.obs.cpu().numpy().reshape(E, N, 204)
        tobs_g = gpu.tobs.cpu().numpy().reshape(E, N, 204)
        for e in range(E):
            n_tie += compare_obs('t=%d e=%d' % (t, e), obs_g[e], obs_o[e], orc.state['pos'][e])
            if d[e]:
                same = (slot_ids(tobs_g[e]) == slot_ids(orc.tobs[e]))[..., None] & np.ones((1, 1, 17), bool)
                gu.assert_close_rel(np.where(same, tobs_g[e].reshape(-1, 12, 17), 0),
                                    np.where(same, orc.tobs[e].reshape(-1, 12, 17), 0), RTOL, 'tobs', atol=1e-12)
    assert n_done >= E
    assert n_tie <= 1e-4 * E * N * 12 * steps  # same tie rule on both sides: only f32-rounding flips of ~1e-16 offsets


@pytest.mark.gpu
def test_cuda_mt_superpool_facade_matches_reference_autoreset():
    import torch
    from hmp2g_b200.multiteam import DcaMultiTeamSuperpoolEnv
    name = '3teams'
    case = load_case(name)
    host = DcaMultiTeamSuperpoolEnv(n_env=1, scfg=_mt_scfg(case['cfg']), seed=case['seed'], env_gid0=case['gid'])
    N = host.env.N
    obs0, info0 = host.reset()
    assert obs0.shape == (1, N, 204) and obs0.dtype == np.float32 and info0[0] == {}
    compare_obs('obs0', obs0[0], case['obs0'], case['s0_pos'])
    host.env.import_state({k: case['s0_' + k][None] for k in STATE_KEYS})
    want_tobs = {int(t): case['tobs'][k] for k, t in enumerate(case['tobs_steps'])}
    last_info = None
    n_done = n_skip = 0
    for t in range(case['actions'].shape[0]):
        a = case['actions'][t]
        obs, rew, done, info = host.step(a[None])
        tag = '%s t=%d' % (name, t)
        assert rew.dtype == np.float64 and done.dtype == bool and info.dtype == object
        assert bool(done[0]) == bool(case['done'][t]), tag
        assert np.array_equal(rew[0], case['reward'][t]), tag
        compare_obs(tag, obs[0], case['obs'][t], case['pos'][t])
        if np.isnan(a).any():
            n_skip += 1
            assert set(info[0]) == set(last_info), tag
        elif done[0]:
            n_done += 1
            assert set(info[0]) == {'team_ranking', 'end_reason', 'obs-echo'}, tag
            assert np.array_equal(info[0]['team_ranking'], case['ranking'][t]) and info[0]['end_reason'] == 'EndGame'
            assert np.array_equal(slot_ids(info[0]['obs-echo']) >= 0, slot_ids(want_tobs[t]) >= 0), tag
        else:
            assert info[0] == {}, tag
        last_info = info[0]
    assert n_done >= 2 and n_skip >= 2


@pytest.mark.gpu
def test_cuda_mt_unity_modes_match_reference():
    """RewardAsUnity / ObsAsUnity (MISSION/dca_multiteam/collective_assault_parallel_run.py:166-186): one reward per
    team (0, then the ranking bonus on done) and agent 0's observation alone, against a reference trajectory recorded
    with both flags on (tests/golden/mt_4x6_unity.npz), through the device and the host surface."""
    import torch
    from hmp2g_b200.multiteam import DcaMultiTeamSuperpoolEnv, DcaMultiTeamVecEnv
    case = load_case('4x6_unity')
    plain = load_case('4x6')
    assert case['cfg']['RewardAsUnity'] and case['cfg']['ObsAsUnity']
    T = len(case['cfg']['N_AGENT_EACH_TEAM'])
    env = DcaMultiTeamVecEnv(1, scfg=_mt_scfg(case['cfg']), seed=case['seed'], env_gid0=case['gid'])
    host = DcaMultiTeamSuperpoolEnv(n_env=1, scfg=_mt_scfg(case['cfg']), seed=case['seed'], env_gid0=case['gid'])
    obs0, _ = host.reset()
    assert obs0.shape == (1, 1, 204)
    for e in (env, host.env):
        e.import_state({k: case['s0_' + k][None] for k in STATE_KEYS})
    want_tobs = {int(t): case['tobs'][k] for k, t in enumerate(case['tobs_steps'])}
    n_done = 0
    for t in range(case['actions'].shape[0]):
        a = case['actions'][t]
        skip = bool(np.isnan(a).any())
        act = torch.from_numpy(np.where(np.isnan(a), 0, a).astype(np.int32)[None]).to(env.device)
        obs, rew, done, rank = env.step(act, skip=torch.full((1,), int(skip), dtype=torch.uint8, device=env.device))
        tag = 'unity t=%d' % t
        assert tuple(obs.shape) == (1, 1, 12, 17) and tuple(rew.shape) == (1, T)
        assert int(done[0]) == int(case['done'][t]), tag
        assert np.array_equal(rew[0].cpu().numpy().astype(np.float64), case['reward'][t]), tag
        # agent 0's rows: the same comparison as the per-agent fixture, restricted to the first agent
        o = obs[0].reshape(1, 204).cpu().numpy()
        ida, idb = slot_ids(o), slot_ids(case['obs'][t])
        assert np.array_equal(ida >= 0, idb >= 0), tag
        same = (ida == idb)[..., None] & np.ones((1, 1, 17), bool)
        gu.assert_close_rel(np.where(same, o.reshape(1, 12, 17), 0), np.where(same, case['obs'][t].reshape(1, 12, 17), 0),
                            RTOL, tag + ' obs', atol=1e-12)
        h_obs, h_rew, h_done, h_info = host.step(a[None])
        assert h_obs.shape == (1, 1, 204) and h_rew.shape == (1, T) and h_rew.dtype == np.float64
        assert np.array_equal(h_rew[0], case['reward'][t]) and bool(h_done[0]) == bool(case['done'][t]), tag
        assert np.array_equal(h_obs[0], o), tag
        if case['done'][t] and not skip:
            n_done += 1
            assert np.array_equal(rank[0].cpu().numpy(), case['ranking'][t]), tag
            assert np.array_equal(h_info[0]['team_ranking'], case['ranking'][t]) and h_info[0]['obs-echo'].shape == (1, 204)
            assert np.array_equal(slot_ids(h_info[0]['obs-echo']) >= 0, slot_ids(want_tobs[t]) >= 0), tag
    assert n_done >= 5
    assert np.array_equal(case['done'], plain['done'])  # the flags change what is handed out, not the simulation
