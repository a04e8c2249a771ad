"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C-ABI
(hmp2g_b200._lib -> libdca_b200.so), against (a) the golden trajectories produced by the reference itself and
(b) the CPU oracle on the same seeded inputs.  Bar (BASELINE.json north_star): integer / index outputs bit-exact
(alive masks, kill counts, done, win, kNN slot ids, nearest-guard index); floats within 1e-5 relative per step over
a 100-step horizon.  /root/reference is NOT needed here.
"""
import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # north_star tolerance for float positions / rewards


def _torch():
    import torch
    assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
    return torch


class _Cfg(object):
    pass


def scenario_from_dict(d):
    """A ScenarioConfig-like class carrying the case's parameters (what JSONC injection would produce)."""
    from hmp2g_b200.config import ScenarioConfig
    c = type('CaseScenarioConfig', (ScenarioConfig,), {})
    for k, v in d.items():
        assert hasattr(ScenarioConfig, k), k
        setattr(c, k, v)
    return c


def make_gpu_env(cfg, n_env, seed, gid0, debug=True):
    from hmp2g_b200.vec_env import DcaVecEnv
    return DcaVecEnv(n_env, scfg=scenario_from_dict(cfg), seed=seed, env_gid0=gid0, debug_outputs=debug)


def compare_state(gpu_state, ref_state, tag, e_gpu=slice(None), rtol=RTOL):
    for k in gu.INT_KEYS:
        a = gpu_state[k][e_gpu].astype(np.int64)
        b = np.asarray(ref_state[k]).astype(np.int64)
        if k in ('obs_counter', 'episode'):
            a, b = a & 0xFFFFFFFF, b & 0xFFFFFFFF
        assert np.array_equal(a, b), (tag, k)
    for k in gu.FLOAT_KEYS:
        gu.assert_close_rel(gpu_state[k][e_gpu], ref_state[k], rtol, '%s %s' % (tag, k), atol=1e-12)


@pytest.mark.parametrize('name', gu.golden_cases())
def test_cuda_matches_reference_golden(name):
    """Replay the reference's recorded trajectory (state0, guard actions, planner forces) on the GPU."""
    torch = _torch()
    case = gu.load_case(name)
    env = make_gpu_env(case['cfg'], 2, case['seed'], case['gid'])  # env 1 is a bystander with another global id
    s0 = gu.state0(case)
    env.import_state({k: np.stack([v, v]) for k, v in s0.items()})
    T = case['actions'].shape[0]
    dev = env.device
    n_bits_equal = n_vals = 0
    for t in range(T):
        a = case['actions'][t]
        skip = np.isnan(a).any()
        act = torch.from_numpy(np.tile(np.where(np.isnan(a), 0, a).astype(np.int32), (2, 1))).to(dev)
        frc = torch.from_numpy(np.tile(case['att_force'][t], (2, 1, 1))).to(dev)
        skp = torch.full((2,), int(skip), dtype=torch.uint8, device=dev)
        env.step(act, frc, skip=skp)
        st = env.export_state()
        tag = '%s t=%d' % (name, t)
        compare_state(st, {k: case[k][t] for k in gu.STATE_KEYS}, tag, e_gpu=0)
        assert int(env.done[0]) == int(case['done'][t]), tag
        if case['done'][t]:
            assert int(env.win[0]) == int(case['win'][t]), tag
        assert np.array_equal(env.reward[0].cpu().numpy().astype(np.float64), case['reward'][t]), tag
        obs = env.obs[0].reshape(env.G, 192).cpu().numpy()
        ref_ids = gu.slot_ids_from_obs(case['obs'][t])
        assert np.array_equal(env.slot_id[0].cpu().numpy().astype(np.int32), ref_ids), tag
        assert np.array_equal(gu.slot_ids_from_obs(obs), ref_ids), tag
        gu.assert_close_rel(obs, case['obs'][t], RTOL, tag + ' obs', atol=1e-12)
        n_bits_equal += int((obs.view(np.uint32) == case['obs'][t].view(np.uint32)).sum())
        n_vals += obs.size
        if not skip:
            assert np.array_equal(env.nearest[0].cpu().numpy(), case['nearest'][t]), tag
        assert np.array_equal(env.alive[0].cpu().numpy(), case['alive'][t]), tag
    for k, t in enumerate(case['tobs_steps']):
        pass  # terminal obs is checked in test_cuda_terminal_obs (needs the value at the done step)
    print('%s: obs float32 bit-identical to the reference in %.4f%% of %d values' % (name, 100.0 * n_bits_equal / n_vals, n_vals))


@pytest.mark.parametrize('name', ['cfgA_50v50', 'small_8v10_fluct'])
def test_cuda_terminal_obs(name):
    torch = _torch()
    case = gu.load_case(name)
    env = make_gpu_env(case['cfg'], 1, case['seed'], case['gid'])
    env.import_state({k: v[None] for k, v in gu.state0(case).items()})
    want = {int(t): case['tobs'][k] for k, t in enumerate(case['tobs_steps'])}
    assert want
    for t in range(case['actions'].shape[0]):
        a = case['actions'][t]
        skip = np.isnan(a).any()
        act = torch.from_numpy(np.where(np.isnan(a), 0, a).astype(np.int32)[None]).to(env.device)
        frc = torch.from_numpy(case['att_force'][t][None]).to(env.device)
        env.step(act, frc, skip=torch.full((1,), int(skip), dtype=torch.uint8, device=env.device))
        if t in want:
            tobs = env.tobs[0].reshape(env.G, 192).cpu().numpy()
            assert np.array_equal(gu.slot_ids_from_obs(tobs), gu.slot_ids_from_obs(want[t])), (name, t)
            gu.assert_close_rel(tobs, want[t], RTOL, 'tobs', atol=1e-12)


@pytest.mark.parametrize('name', gu.golden_cases())
def test_cuda_reset_matches_reference(name):
    """In-kernel Philox + Box-Muller reset against the reference's reset under the same keyed draws."""
    _torch()
    case = gu.load_case(name)
    env = make_gpu_env(case['cfg'], 1, case['seed'], case['gid'])
    env.reset()
    st = env.export_state()
    for k in gu.INT_KEYS:
        if k == 'just_died':
            continue
        assert np.array_equal(st[k][0].astype(np.int64) & 0xFFFFFFFF, case['s0_' + k].astype(np.int64) & 0xFFFFFFFF), k
    for k in gu.FLOAT_KEYS:
        gu.assert_close_rel(st[k][0], case['s0_' + k], 1e-9, 'reset ' + k, atol=1e-12)
    obs = env.obs[0].reshape(env.G, 192).cpu().numpy()
    assert np.array_equal(gu.slot_ids_from_obs(obs), gu.slot_ids_from_obs(case['obs0']))
    gu.assert_close_rel(obs, case['obs0'], 1e-6, 'reset obs', atol=1e-7)


def _run_against_oracle(cfg, E, steps, seed, gid0, nthreads=8, use_stream_every=0, resync=True):
    """GPU and oracle step side by side on identical seeded inputs; returns summary counters."""
    torch = _torch()
    import dca_rng
    from dca_oracle import OracleVecEnv
    orc = OracleVecEnv(cfg, E, seed=seed, env_gid0=gid0, nthreads=nthreads)
    env = make_gpu_env(cfg, E, seed, gid0)
    orc.reset()
    env.reset()
    st = env.export_state()
    compare_state(st, orc.state, 'reset', rtol=1e-9)
    env.import_state(orc.state)  # start both from bit-identical states
    rng = np.random.RandomState(seed + 17)
    n_done = n_kill = 0
    for t in range(steps):
        act = rng.randint(0, 7, size=(E, orc.G)).astype(np.int32)
        skip = np.zeros((E,), np.uint8)
        if t % 7 == 3:
            skip = ((orc.done != 0) & (rng.rand(E) < 0.5)).astype(np.uint8)  # pause some just-reset envs
        force = orc.plan_centroid()
        gforce = env.plan_centroid().cpu().numpy()
        gu.assert_close_rel(gforce, force, 1e-12, 'centroid planner t=%d' % t, atol=1e-13)
        words = None
        if use_stream_every and t % use_stream_every == 0:
            words = dca_rng.obs_words_all(seed, gid0 + np.arange(E),
                                          (orc.state['obs_counter'].astype(np.int64) + 1) & 0xFFFFFFFF, orc.G)
        act_f = act.astype(np.float64)
        act_f[skip != 0] = np.nan
        orc.step(act_f, force, rng_words=words)
        env.step(torch.from_numpy(act).to(env.device), torch.from_numpy(force).to(env.device),
                 skip=torch.from_numpy(skip).to(env.device),
                 rng_words=None if words is None else torch.from_numpy(words.view(np.int32)).to(env.device))
        tag = 't=%d' % t
        st = env.export_state()
        compare_state(st, orc.state, tag)
        assert np.array_equal(env.done.cpu().numpy(), orc.done), tag
        d = orc.done != 0
        assert np.array_equal(env.win.cpu().numpy()[d], orc.win[d]), tag
        assert np.array_equal(env.reward.cpu().numpy().astype(np.float64), orc.reward), tag
        assert np.array_equal(env.slot_id.cpu().numpy(), orc.slot_id), tag
        ns = skip == 0
        assert np.array_equal(env.nearest.cpu().numpy()[ns], orc.nearest[ns]), tag
        obs = env.obs.reshape(E, orc.G, 192).cpu().numpy()
        gu.assert_close_rel(obs, orc.obs, RTOL, tag + ' obs', atol=1e-12)
        if d.any():
            tob = env.tobs.reshape(E, orc.G, 192).cpu().numpy()
            dd = d & ns
            assert np.array_equal(gu.slot_ids_from_obs(tob[dd]), gu.slot_ids_from_obs(orc.tobs[dd])), tag
            gu.assert_close_rel(tob[dd], orc.tobs[dd], RTOL, tag + ' tobs', atol=1e-12)
        n_done += int((d & ns).sum())
        n_kill += int(orc.state['just_died'].sum())
        if resync:
            env.import_state(orc.state)  # keep ulp-level libm differences from accumulating into index flips
    return dict(n_done=n_done, n_kill=n_kill)


def test_cuda_vs_oracle_50v50_100_steps_no_resync():
    """The north_star statement verbatim: identical initial states + action streams, 100 steps, no re-sync."""
    cfg = dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True,
               terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180)
    r = _run_against_oracle(cfg, 48, 100, seed=5, gid0=1000, resync=False)
    assert r['n_kill'] > 100


def test_cuda_vs_oracle_short_episodes_autoreset_skip_stream():
    cfg = dict(num_guards=12, num_attackers=14, random_jam_prob=0.2, introduce_terrain=True,
               terrain_parameters=[0.1, 0.3], MaxEpisodeStep=25, init_distance=2, size=3.0,
               enable_attacker_fluctuation=True, attacker_fluctuation=5, REWARD_DEBUG=True, REWARD_DEBUG_Value=3)
    r = _run_against_oracle(cfg, 96, 120, seed=11, gid0=2 ** 31 - 40, use_stream_every=3)
    assert r['n_done'] > 96 * 3


def test_cuda_vs_oracle_100v100():
    cfg = dict(num_guards=100, num_attackers=100, random_jam_prob=0.05, introduce_terrain=True,
               terrain_parameters=[0.05, 0.2], MaxEpisodeStep=60)
    _run_against_oracle(cfg, 16, 70, seed=3, gid0=0)


def test_cuda_vs_oracle_ragged_150v131():
    """Largest sizes the reference admits (0..150 per side) with an odd N: blob padding / 256-thread path."""
    cfg = dict(num_guards=150, num_attackers=131, random_jam_prob=0.5, introduce_terrain=False, MaxEpisodeStep=30)
    _run_against_oracle(cfg, 6, 35, seed=8, gid0=77)


def test_cuda_full_size_4096_envs_vs_oracle():
    """BASELINE config 2 at full size: 4096 envs x 50v50; every env checked against the oracle for 40 steps."""
    cfg = dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True,
               terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180)
    r = _run_against_oracle(cfg, 4096, 40, seed=2024, gid0=0)
    assert r['n_kill'] > 4096


def test_cuda_full_size_invariants_long_run():
    """Size-independent properties at 4096 envs over whole episodes (no oracle): every hit has a shooter and a
    target (sum numHit == sum numWasHit), alive mask == 1 - ever-hit, obs rows equal the entity records they index,
    dead guards observe zeros, done envs come back fully alive at time_step 0."""
    torch = _torch()
    cfg = dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True,
               terrain_parameters=[0.05, 0.2], MaxEpisodeStep=60)
    E = 4096
    env = make_gpu_env(cfg, E, 9, 0)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(1)
    n_done = 0
    for t in range(130):
        act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
        env.step(act, None)
        if t % 10 == 9 or t in (59, 60, 119, 120):
            st = env.export_state()
            assert (st['num_hit'].sum(1) == st['num_was_hit'].sum(1)).all()
            assert np.array_equal(st['alive'] != 0, st['num_was_hit'] == 0) or cfg.get('enable_attacker_fluctuation')
            assert (st['n_alive_guards'] == st['alive'][:, :env.G].sum(1)).all()
            assert (st['n_alive_attackers'] == st['alive'][:, env.G:].sum(1)).all()
            obs = env.obs.cpu().numpy()
            sid = env.slot_id.cpu().numpy().astype(np.int64)
            assert np.array_equal(gu.slot_ids_from_obs(obs.reshape(E, env.G, 192)), sid)
            dead_g = st['alive'][:, :env.G] == 0
            assert (obs[dead_g] == 0).all()
            # gathered rows are exactly the float32 entity records
            e_idx, g_idx, r_idx = np.nonzero(sid >= 0)
            ent = sid[e_idx, g_idx, r_idx]
            rows = obs[e_idx, g_idx, r_idx]
            assert np.array_equal(rows[:, 1], st['pos'][e_idx, ent, 0].astype(np.float32))
            assert np.array_equal(rows[:, 2], st['pos'][e_idx, ent, 1].astype(np.float32))
            assert np.array_equal(rows[:, 7], st['terrain'][e_idx, ent].astype(np.float32))
            assert (sid[:, :, 0][~dead_g] == np.broadcast_to(np.arange(env.G), (E, env.G))[~dead_g]).all()
            d = env.done.cpu().numpy() != 0
            n_done += int(d.sum())
            if d.any():
                assert (st['time_step'][d] == 0).all() and (st['alive'][d] == 1).all()
    assert n_done >= E  # every env hit the 60-step limit at least once at t=59 / t=119


def test_cuda_gae_matches_reference_and_oracle():
    torch = _torch()
    from dca_oracle import gae as orc_gae
    from hmp2g_b200.gae import gae_return
    z = np.load(gu.GOLDEN_DIR + '/gae_conc.npz')
    n = len([k for k in z.files if k.endswith('_reward')])
    for i in range(n):
        r, v, d = z['c%d_reward' % i], z['c%d_value' % i], z['c%d_dead' % i]
        dev = 'cuda'
        ret, threat, pushed = gae_return(torch.from_numpy(r[:, None].astype(np.float32)).to(dev).contiguous(),
                                         torch.from_numpy(v[:, None]).to(dev).contiguous(),
                                         torch.from_numpy(d[:, None]).to(dev), want_pushed=True)
        assert np.array_equal(pushed[:, 0].cpu().numpy().astype(np.float64), z['c%d_reward_pushed' % i])
        assert np.array_equal(threat[:, 0].cpu().numpy(), z['c%d_threat' % i])
        assert np.array_equal(ret[:, 0].cpu().numpy().view(np.uint32), z['c%d_return' % i].view(np.uint32)), i
    # ragged trajectory lengths at rollout-buffer size, vs the oracle
    rng = np.random.RandomState(0)
    T, E, G = 180, 256, 50
    r = rng.choice([0.0, 1.0, -0.5, 3.5], size=(T, E, G))
    v = rng.randn(T, E, G).astype(np.float32)
    death = rng.randint(1, T + 40, size=(E, G))
    d = np.arange(T)[:, None, None] >= death[None]
    tl = rng.randint(0, T + 1, size=E).astype(np.int32)
    tl[:3] = (0, 1, T)
    o_ret, o_thr, o_push = orc_gae(r, v, d, traj_len=tl, threat_limit=8)
    ret, thr, push = gae_return(torch.from_numpy(r.astype(np.float32)).cuda(), torch.from_numpy(v).cuda(),
                                torch.from_numpy(d).cuda(), traj_len=torch.from_numpy(tl).cuda(), threat_limit=8,
                                want_pushed=True)
    assert np.array_equal(ret.cpu().numpy().view(np.uint32), o_ret.view(np.uint32))
    assert np.array_equal(thr.cpu().numpy(), o_thr)
    assert np.array_equal(push.cpu().numpy().astype(np.float64), o_push)
    # gamma_in_reward_forwarding (conc_4hist/trajectory.py:57-63): the reference's own fixtures, then ragged vs the oracle
    z = np.load(gu.GOLDEN_DIR + '/gae_conc4hist_fwd.npz')
    for tag in 'ab':
        gf = float(z[tag + '_gamma_fwd'])
        for i in range(3):
            r_, v_, d_ = (z['%s_c%d_%s' % (tag, i, k)] for k in ('reward', 'value', 'dead'))
            ret_, thr_, push_ = gae_return(torch.from_numpy(r_[:, None].astype(np.float32)).cuda().contiguous(),
                                           torch.from_numpy(v_[:, None]).cuda().contiguous(),
                                           torch.from_numpy(d_[:, None]).cuda(), want_pushed=True, gamma_fwd=gf)
            want_push = z['%s_c%d_reward_pushed' % (tag, i)]
            # rewards enter as float32 here (0.3 and -1.7 are not float32 numbers): compare at float32 resolution
            np.testing.assert_allclose(push_[:, 0].cpu().numpy(), want_push, rtol=2e-7, atol=1e-7)
            assert np.array_equal(thr_[:, 0].cpu().numpy(), z['%s_c%d_threat' % (tag, i)])
            np.testing.assert_allclose(ret_[:, 0].cpu().numpy(), z['%s_c%d_return' % (tag, i)], rtol=1e-6, atol=1e-6)
    r32 = r.astype(np.float32)
    o_ret, o_thr, o_push = orc_gae(r32.astype(np.float64), v, d, traj_len=tl, threat_limit=11, gamma_fwd=0.97)
    ret, thr, push = gae_return(torch.from_numpy(r32).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(d).cuda(),
                                traj_len=torch.from_numpy(tl).cuda(), threat_limit=11, want_pushed=True, gamma_fwd=0.97)
    assert np.array_equal(ret.cpu().numpy().view(np.uint32), o_ret.view(np.uint32))  # same f64 operations, one rounding
    assert np.array_equal(thr.cpu().numpy(), o_thr)
    assert np.array_equal(push.cpu().numpy(), o_push.astype(np.float32))


def test_cuda_empty_rows_nan_option_and_device_rollout():
    """(1) DCA_OPT_EMPTY_ROWS_NAN hands over exactly what ShellEnvWrapper makes of the env's observation
    (ALGORITHM/conc/shell_env.py:72-75: all-zero rows -> NaN).  (2) DeviceRollout + dca_gae reproduce
    trajectory.finalize (restated by the oracle's gae()) on the rewards / dead masks of a rollout with the native planner."""
    torch = _torch()
    from dca_oracle import gae as oracle_gae
    from hmp2g_b200.rollout import DeviceRollout
    from hmp2g_b200.vec_env import DcaVecEnv
    cfg = dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True,
               terrain_parameters=[0.05, 0.2], MaxEpisodeStep=50, init_distance=2)
    E, T = 64, 50
    plain = make_gpu_env(cfg, E, seed=9, gid0=0, debug=False)
    nan_env = DcaVecEnv(E, scfg=scenario_from_dict(cfg), seed=9, env_gid0=0, empty_rows_nan=True)
    plain.reset(); nan_env.reset()
    roll = DeviceRollout(nan_env, horizon=T)
    gen = torch.Generator(device=plain.device); gen.manual_seed(4)
    rec_r, rec_v, rec_d = [], [], []
    finished_any = torch.zeros((E,), dtype=torch.bool, device=plain.device)
    for t in range(T):
        act = torch.randint(0, 7, (E, 50), generator=gen, device=plain.device, dtype=torch.int32)
        alive_before = roll.alive_guards().clone()
        value = torch.rand((E, 50), generator=gen, device=plain.device)
        force = plain.plan().clone()
        plain.step(act, force)
        nan_env.step(act, force)
        o0, o1 = plain.obs.view(E, 50, 12, 16), nan_env.obs.view(E, 50, 12, 16)
        want = torch.where((o0 == 0).all(-1, keepdim=True), torch.full_like(o0, float('nan')), o0)
        assert torch.equal(torch.isnan(o1), torch.isnan(want)) and torch.equal(torch.nan_to_num(o1), torch.nan_to_num(want)), t
        assert torch.equal(plain.reward, nan_env.reward) and torch.equal(plain.done, nan_env.done)
        if t == 0 or finished_any.any():
            pass
        finished_any |= roll.add(alive_before, value, nan_env.reward, nan_env.done)
        rec_r.append(nan_env.reward.cpu().numpy().copy()); rec_v.append(value.cpu().numpy().copy())
        rec_d.append((alive_before == 0).cpu().numpy().copy())
        if bool(nan_env.done.any()):
            break
    # every env finalized as a (possibly truncated) trajectory of the steps collected so far
    n = len(rec_r)
    ret, threat, adv, valid = roll.finalize()
    r = np.stack(rec_r).astype(np.float64); v = np.stack(rec_v); d = np.stack(rec_d)
    want_ret, want_thr, _ = oracle_gae(r, v, d, traj_len=None)
    np.testing.assert_allclose(ret[:n].cpu().numpy(), want_ret, rtol=1e-6, atol=1e-6)
    assert np.array_equal(threat[:n].cpu().numpy(), want_thr)
    assert bool(valid[:n].all()) and not bool(valid[n:].any())
    a = adv[:n].cpu().numpy()
    assert abs(a.mean()) < 1e-3 and abs(a.std() - 1) < 1e-2
    assert int(roll.length.sum()) == 0


@pytest.mark.parametrize('name', ['cfgA_50v50', 'small_8v10_fluct', 'small_6v10_nored'])
def test_cuda_superpool_facade_matches_reference_autoreset(name):
    """DcaSuperpoolEnv (the SuperpoolEnv / EnvAutoReset surface, UTIL/shm_env.py:33-97, :126-168) against the
    reference's own EnvAutoReset trajectory: host float64 actions with NaN (ENV-PAUSE) rows in, numpy obs / float64
    reward / bool done / info dicts out, 'obs-echo' + post-reset obs on done, cached result on a paused step."""
    torch = _torch()
    from hmp2g_b200.vec_env import DcaSuperpoolEnv
    case = gu.load_case(name)
    G = int(case['cfg']['num_guards'])
    t_now = [0]

    def recorded_planner(env):
        return torch.from_numpy(case['att_force'][t_now[0]][None]).to(env.device)

    host = DcaSuperpoolEnv(n_env=1, scfg=scenario_from_dict(case['cfg']), seed=case['seed'], env_gid0=case['gid'],
                           planner=recorded_planner)
    assert host.num_envs == 1 and host.get_space() == {'obs_space': [], 'act_space': []}
    obs0, info0 = host.reset()
    assert obs0.shape == (1, G, 192) and obs0.dtype == np.float32 and info0.shape == (1,)
    assert np.array_equal(gu.slot_ids_from_obs(obs0[0]), gu.slot_ids_from_obs(case['obs0']))
    gu.assert_close_rel(obs0[0], case['obs0'], 1e-6, name + ' obs0', atol=1e-7)
    obs_again, _ = host.reset()  # later manual resets return the cached post-reset result (shm_env.py:82-93)
    assert np.array_equal(obs_again, obs0)
    # continue from the reference's own post-reset state (the in-kernel Box-Muller reset equals it to ~1e-9, see
    # test_cuda_reset_matches_reference; the trajectory comparison below should not depend on that)
    host.env.import_state({k: v[None] for k, v in gu.state0(case).items()})
    want_tobs = {int(t): case['tobs'][k] for k, t in enumerate(case['tobs_steps'])}
    n_done = n_skip = 0
    last_info = None
    for t in range(case['actions'].shape[0]):
        t_now[0] = t
        a = case['actions'][t]
        skip = bool(np.isnan(a).any())
        obs, rew, done, info = host.step(a[None])
        tag = '%s t=%d' % (name, t)
        assert obs.dtype == np.float32 and rew.dtype == np.float64 and done.dtype == bool and info.dtype == object
        assert bool(done[0]) == bool(case['done'][t]), tag
        assert np.array_equal(rew[0], case['reward'][t]), tag
        assert np.array_equal(gu.slot_ids_from_obs(obs[0]), gu.slot_ids_from_obs(case['obs'][t])), tag
        gu.assert_close_rel(obs[0], case['obs'][t], RTOL, tag + ' obs', atol=1e-12)
        if skip:  # the cached previous (ob, reward, done, info), shm_env.py:37-41
            n_skip += 1
            assert info[0] is last_info or info[0] == last_info or set(info[0]) == set(last_info), tag
        elif done[0]:
            n_done += 1
            assert set(info[0]) == {'win', 'obs-echo', 'others'}, tag
            assert bool(info[0]['win']) == bool(case['win'][t]), tag
            echo = np.asarray(info[0]['obs-echo'])
            assert echo.shape == (G, 192)
            assert np.array_equal(gu.slot_ids_from_obs(echo), gu.slot_ids_from_obs(want_tobs[t])), tag
            gu.assert_close_rel(echo, want_tobs[t], RTOL, tag + ' obs-echo', atol=1e-12)
        else:
            assert info[0] == {}, tag
        last_info = info[0]
    assert n_done >= 1 and n_skip >= 1


def test_cuda_superpool_plan_ahead_equals_plan_in_step():
    """The façade computes the scripted attackers' plan for step t+1 under the device->host copy of step t; the
    trajectory must equal the one of a DcaVecEnv that plans inside step()."""
    torch = _torch()
    from hmp2g_b200.vec_env import DcaSuperpoolEnv, DcaVecEnv
    cfg = dict(num_guards=20, num_attackers=30, MaxEpisodeStep=25)
    E = 64
    host = DcaSuperpoolEnv(n_env=E, scfg=scenario_from_dict(cfg), seed=5, env_gid0=100)
    env = DcaVecEnv(E, scfg=scenario_from_dict(cfg), seed=5, env_gid0=100)
    o_h, _ = host.reset()
    env.reset()
    assert np.array_equal(o_h.reshape(-1), env.obs.cpu().numpy().reshape(-1))
    rng = np.random.RandomState(0)
    n_done = 0
    for t in range(60):
        a = rng.randint(0, 7, size=(E, 20))
        o_h, r_h, d_h, info = host.step(a.astype(np.float64))
        env.step(torch.from_numpy(a.astype(np.int32)).to(env.device))
        assert np.array_equal(o_h.view(np.uint32).reshape(-1), env.obs.cpu().numpy().view(np.uint32).reshape(-1)), t
        assert np.array_equal(r_h, env.reward.cpu().numpy().astype(np.float64)), t
        assert np.array_equal(d_h, env.done.cpu().numpy().astype(bool)), t
        n_done += int(d_h.sum())
    assert n_done >= E


@pytest.mark.parametrize('planner', ['centroid', 'scripted'])
def test_cuda_step_overlapping_its_planner_is_bit_identical(planner):
    """DCA_OPT_OVERLAP_PLANNER (the step kernel launched as a programmatic dependent of the planner kernel, ordering
    itself behind it only before it reads the forces) must not change a single bit: full-size 50v50 batch, both planner
    kernels, outputs and exported state compared with a strictly serial run."""
    torch = _torch()
    from hmp2g_b200.vec_env import DcaVecEnv
    cfg = dict(num_guards=50, num_attackers=50, MaxEpisodeStep=40, random_jam_prob=0.05, introduce_terrain=True)
    E = 4096
    envs = [DcaVecEnv(E, scfg=scenario_from_dict(cfg), seed=11, env_gid0=0, overlap_planner=o) for o in (True, False)]
    for env in envs:
        env.reset()
    gen = torch.Generator(device=envs[0].device)
    gen.manual_seed(3)
    n_done = 0
    for t in range(90 if planner == 'centroid' else 50):
        a = torch.randint(0, 7, (E, 50), generator=gen, device=envs[0].device, dtype=torch.int32)
        for env in envs:
            force = env.plan_centroid() if planner == 'centroid' else env.plan()
            env.step(a, force)
        for name in ('obs', 'tobs', 'reward', 'done', 'win', 'alive'):
            assert torch.equal(getattr(envs[0], name), getattr(envs[1], name)), (t, name)
        n_done += int(envs[0].done.sum().item())
    assert n_done >= E
    s0, s1 = envs[0].export_state(), envs[1].export_state()
    for k in s0:
        assert np.array_equal(np.asarray(s0[k]).view(np.uint8), np.asarray(s1[k]).view(np.uint8)), k


@pytest.mark.parametrize('planner', ['centroid', 'scripted'])
def test_cuda_pipelined_steps_are_bit_identical(planner):
    """DCA_OPT_PIPELINE_STEPS (plan / step / plan / step issued back to back: every env is handed from launch to launch
    through its own epoch word, the launches overlap at their edges) must not change a single bit: 4096 envs (3.5 waves
    of step CTAs, so launch t+1 really starts inside launch t), bursts of 1..40 launches without anything between
    them, time-limit resets inside, compared with a strictly serial run after every burst."""
    torch = _torch()
    from hmp2g_b200.vec_env import DcaVecEnv
    cfg = dict(num_guards=50, num_attackers=50, MaxEpisodeStep=40, random_jam_prob=0.05, introduce_terrain=True)
    E = 4096
    piped = DcaVecEnv(E, scfg=scenario_from_dict(cfg), seed=11, env_gid0=0, overlap_planner=True, pipeline_steps=True)
    serial = DcaVecEnv(E, scfg=scenario_from_dict(cfg), seed=11, env_gid0=0, overlap_planner=False, pipeline_steps=False)
    for env in (piped, serial):
        env.reset()
    gen = torch.Generator(device=piped.device)
    gen.manual_seed(3)
    pool = [torch.randint(0, 7, (E, 50), generator=gen, device=piped.device, dtype=torch.int32) for _ in range(16)]
    t = 0
    n_done = 0
    # t = 40: every env is done; the long bursts at the end: any env handed over too early would fork for good
    for burst in ([1, 2, 3, 7, 27, 5, 40, 600] if planner == 'centroid' else [1, 2, 7, 30, 12, 150]):
        for env in (piped, serial):
            for k in range(burst):   # nothing but plan + step on the stream inside a burst
                force = env.plan_centroid() if planner == 'centroid' else env.plan()
                env.step(pool[(t + k) % 16], force)
        t += burst
        if burst == 7:  # a paused step in between (pause mask: an ordinary launch that restarts the pipeline)
            skip = (torch.arange(E, device=piped.device) % 3 == 0).to(torch.uint8)
            for env in (piped, serial):
                force = env.plan_centroid() if planner == 'centroid' else env.plan()
                env.step(pool[t % 16], force, skip=skip)
        for name in ('obs', 'tobs', 'reward', 'done', 'win', 'alive'):
            assert torch.equal(getattr(piped, name), getattr(serial, name)), (t, name)
        s0, s1 = piped.export_state(), serial.export_state()
        for k in s0:
            assert np.array_equal(np.asarray(s0[k]).view(np.uint8), np.asarray(s1[k]).view(np.uint8)), (t, k)
        n_done += int(serial.done.sum().item())
    assert t >= 47 and n_done >= E // 4   # (at t = 40 the envs paused once are the ones that have just finished)


@pytest.mark.parametrize('groups', [2, 4])
def test_cuda_env_groups_are_bit_identical_to_one_env(groups):
    """DcaEnvGroups: E envs as independent groups on their own streams (contiguous global env ids) must produce what one
    DcaVecEnv(E) produces, bit for bit -- joined every step (sync=True), and in bursts of unjoined steps (sync=False,
    the groups drifting apart by whole launches)."""
    torch = _torch()
    from hmp2g_b200.vec_env import DcaEnvGroups, DcaVecEnv
    cfg = dict(num_guards=50, num_attackers=50, MaxEpisodeStep=40, random_jam_prob=0.05, introduce_terrain=True)
    E = 2048
    one = DcaVecEnv(E, scfg=scenario_from_dict(cfg), seed=11, env_gid0=64)
    grp = DcaEnvGroups(E, groups=groups, scfg=scenario_from_dict(cfg), seed=11, env_gid0=64)
    one.reset()
    grp.reset()
    gen = torch.Generator(device=one.device)
    gen.manual_seed(3)
    acts = [torch.randint(0, 7, (E, 50), generator=gen, device=one.device, dtype=torch.int32) for _ in range(8)]
    torch.cuda.synchronize()

    def same(t):
        for name in ('obs', 'tobs', 'reward', 'done', 'win', 'alive'):
            assert torch.equal(getattr(one, name), getattr(grp, name)), (t, name)

    same(-1)
    n_done = 0
    for t in range(45):   # joined steps, scripted planner; episodes end and autoreset
        one.step(acts[t % 8], one.plan())
        grp.step(acts[t % 8], planner='scripted')
        same(t)
        n_done += int(one.done.sum().item())
    assert n_done >= E
    for burst in (3, 17, 40):   # unjoined bursts, centroid stand-in
        for t in range(burst):
            one.step(acts[t % 8], one.plan_centroid())
        grp.fork()
        for t in range(burst):
            grp.step(acts[t % 8], planner='centroid', sync=False)
        grp.join()
        same(burst)
    # ENV-PAUSE mask and injected attacker forces are handed to the groups as slices
    force = torch.randn((E, 50, 2), generator=gen, device=one.device, dtype=torch.float64)
    for t in range(6):
        skip = (torch.rand((E,), generator=gen, device=one.device) < 0.3).to(torch.uint8)
        one.step(acts[t % 8], force, skip=skip)
        grp.step(acts[t % 8], att_force=force, skip=skip)
        same(100 + t)
    grp.observe()
    one.observe()
    same(200)
    assert grp.kernel_launches() >= groups * (1 + 2 * (45 + 60))
    grp.close()


def test_cuda_gae_policy_resonance_levels_match_reference():
    """gae_return_levels against the reference's gae_finalize_return_pr fixtures (ALGORITHM/common/traj_gae.py:135-169):
    returns of every value level bit-exact, threat (SAFE_LIMIT 8) and pushed reward exact."""
    torch = _torch()
    from hmp2g_b200.gae import gae_return_levels
    z = np.load(gu.GOLDEN_DIR + '/gae_common_pr.npz')
    for i in range(3):
        r, lv, d = z['c%d_reward' % i], z['c%d_levels' % i], z['c%d_dead' % i]
        ret, thr, push = gae_return_levels(torch.from_numpy(r[:, None].astype(np.float32)).cuda().contiguous(),
                                           torch.from_numpy(lv[:, None]).cuda().contiguous(),
                                           torch.from_numpy(d[:, None]).cuda(), threat_limit=8)
        assert np.array_equal(push[:, 0].cpu().numpy().astype(np.float64), z['c%d_reward_pushed' % i])
        assert np.array_equal(thr[:, 0].cpu().numpy(), z['c%d_threat' % i])
        assert np.array_equal(ret[:, 0].cpu().numpy().view(np.uint32), z['c%d_return_levels' % i].view(np.uint32)), i


def test_superpool_reference_obs_dtype_option():
    """reference_obs_dtype=True: float32 observations while every guard is alive, float64 (same values) once one is dead
    -- what np.stack over the reference's per-env observations yields."""
    torch = _torch()
    from hmp2g_b200.vec_env import DcaSuperpoolEnv
    cfg = dict(num_guards=8, num_attackers=10, MaxEpisodeStep=60, init_distance=1.0, size=2.0)
    a = DcaSuperpoolEnv(n_env=4, scfg=scenario_from_dict(cfg), seed=3, reference_obs_dtype=True)
    b = DcaSuperpoolEnv(n_env=4, scfg=scenario_from_dict(cfg), seed=3)
    oa, _ = a.reset(); ob, _ = b.reset()
    assert oa.dtype == np.float32
    rng = np.random.RandomState(0)
    seen64 = False
    for t in range(60):
        act = rng.randint(0, 7, size=(4, 8)).astype(np.float64)
        oa, ra, da, _ = a.step(act); ob, rb, db, _ = b.step(act)
        assert ob.dtype == np.float32 and np.array_equal(oa.astype(np.float32), ob) and np.array_equal(ra, rb)
        dead = (ob[:, :, 0] == 0).any()
        assert oa.dtype == (np.float64 if dead else np.float32), t
        seen64 |= bool(dead)
    assert seen64


def test_cuda_advantage_normalisation_matches_numpy_formula():
    """adv = ret - value; (adv - adv.mean()) / (adv.std() + 1e-5) as ALGORITHM/conc/ppo.py:53-55 evaluates it on the
    pooled float32 numpy batch (population std); tolerance 1e-5 (the reference sums in float32, the kernel in fp64)."""
    torch = _torch()
    from hmp2g_b200.gae import normalized_advantage
    rng = np.random.RandomState(3)
    for shape in [(180, 64, 50), (7, 3, 5), (1, 1, 2)]:
        ret = (rng.randn(*shape) * 3 + 1).astype(np.float32)
        val = rng.randn(*shape).astype(np.float32)
        adv = ret - val
        want = (adv - adv.mean()) / (adv.std() + 1e-5)
        tr, tv = torch.from_numpy(ret).cuda(), torch.from_numpy(val).cuda()
        got = normalized_advantage(tr, tv)
        assert np.allclose(got.cpu().numpy(), want, rtol=1e-5, atol=1e-5), shape
        again = normalized_advantage(tr, tv)
        assert torch.equal(got, again)  # fixed summation order: bit-reproducible
        # ragged batch: one flag per (t, env) row, and the same mask given elementwise
        rows = rng.rand(*shape[:2]) < 0.6
        rows[0, 0] = True
        sel = adv[rows]
        want_m = (sel - sel.mean()) / (sel.std() + 1e-5)
        got_rows = normalized_advantage(tr, tv, valid=torch.from_numpy(rows).cuda())
        got_elem = normalized_advantage(tr, tv, valid=torch.from_numpy(np.broadcast_to(rows[..., None], shape).copy()).cuda())
        assert torch.equal(got_rows, got_elem)
        g = got_rows.cpu().numpy()
        assert np.allclose(g[rows], want_m, rtol=1e-5, atol=1e-5) and (g[~rows] == 0).all()
