"""CPU: pin oracle/dca_oracle.c against tests/golden (outputs of the reference itself, executed under the
keyed RNG stream by oracle/make_golden.py).  Integer outputs bit-exact; floats 1e-5 relative (BASELINE.json)."""
import numpy as np
import pytest

import golden_util as gu
from dca_oracle import OracleVecEnv, gae

CASES = gu.golden_cases()


def _replay(case, n_copies=1, nthreads=1, use_stream=False):
    import dca_rng
    env = OracleVecEnv(case['cfg'], n_copies, seed=case['seed'], env_gid0=case['gid'], nthreads=nthreads)
    s0 = gu.state0(case)
    for e in range(n_copies):
        env.import_state(e, s0)
    T = case['actions'].shape[0]
    outs = []
    for t in range(T):
        words = None
        if use_stream:
            words = dca_rng.obs_words_all(case['seed'], [case['gid']] * n_copies,
                                          (env.state['obs_counter'].astype(np.int64) + 1) & 0xFFFFFFFF, env.G)
        env.step(np.tile(case['actions'][t], (n_copies, 1)), np.tile(case['att_force'][t], (n_copies, 1, 1)),
                 rng_words=words)
        outs.append(dict(obs=env.obs.copy(), reward=env.reward.copy(), done=env.done.copy(), win=env.win.copy(),
                         tobs=env.tobs.copy(), slot_id=env.slot_id.copy(), nearest=env.nearest.copy(),
                         state={k: v.copy() for k, v in env.state.items()}))
    return outs


@pytest.mark.parametrize('name', CASES)
def test_oracle_matches_reference_trajectory(name):
    case = gu.load_case(name)
    outs = _replay(case)
    T = len(outs)
    n_kill = 0
    for t in range(T):
        o = outs[t]
        skipped = np.isnan(case['actions'][t]).any()
        for k in gu.INT_KEYS:
            assert np.array_equal(o['state'][k][0].astype(np.int64), case[k][t].astype(np.int64)), (name, t, k)
        for k in gu.FLOAT_KEYS:
            gu.assert_close_rel(o['state'][k][0], case[k][t], 1e-5, '%s t=%d %s' % (name, t, k), atol=1e-12)
        assert o['done'][0] == case['done'][t], (name, t)
        if case['done'][t]:
            assert o['win'][0] == case['win'][t], (name, t)
        assert np.array_equal(o['reward'][0], case['reward'][t]), (name, t)
        # slot ids: bit-exact (all gathered rows are at finite distance by construction)
        ref_ids = gu.slot_ids_from_obs(case['obs'][t])
        assert np.array_equal(o['slot_id'][0], ref_ids), (name, t)
        assert np.array_equal(gu.slot_ids_from_obs(o['obs'][0]), ref_ids), (name, t)
        gu.assert_close_rel(o['obs'][0], case['obs'][t], 1e-5, '%s t=%d obs' % (name, t), atol=1e-12)
        if not skipped:
            assert np.array_equal(o['nearest'][0], case['nearest'][t]), (name, t)
        n_kill += int(case['just_died'][t].sum())
    for k, t in enumerate(case['tobs_steps']):
        assert np.array_equal(gu.slot_ids_from_obs(outs[t]['tobs'][0]), gu.slot_ids_from_obs(case['tobs'][k]))
        gu.assert_close_rel(outs[t]['tobs'][0], case['tobs'][k], 1e-5, 'tobs', atol=1e-12)


@pytest.mark.parametrize('name', CASES)
def test_oracle_reset_matches_reference(name):
    case = gu.load_case(name)
    env = OracleVecEnv(case['cfg'], 1, seed=case['seed'], env_gid0=case['gid'])
    env.reset()
    for k in gu.INT_KEYS:
        if k == 'just_died':
            continue
        assert np.array_equal(env.state[k][0].astype(np.int64), case['s0_' + k].astype(np.int64)), k
    for k in gu.FLOAT_KEYS:
        gu.assert_close_rel(env.state[k][0], case['s0_' + k], 1e-9, 'reset ' + k, atol=1e-12)
    assert np.array_equal(gu.slot_ids_from_obs(env.obs[0]), gu.slot_ids_from_obs(case['obs0']))
    gu.assert_close_rel(env.obs[0], case['obs0'], 1e-6, 'reset obs', atol=1e-7)


def test_oracle_stream_equals_philox_and_threads():
    """rng_words stream (what the C-ABI's rng_stream takes) == in-oracle Philox; OpenMP replicas identical."""
    case = gu.load_case('small_8v10_fluct')
    a = _replay(case, n_copies=1)
    b = _replay(case, n_copies=3, nthreads=3, use_stream=True)
    for t in range(len(a)):
        if case['done'][t]:
            break  # after the autoreset the replicas (different global env ids) diverge by design
        for e in range(3):
            assert np.array_equal(a[t]['obs'][0], b[t]['obs'][e])
            assert np.array_equal(a[t]['state']['pos'][0], b[t]['state']['pos'][e])


def test_gae_matches_reference():
    z = np.load(gu.GOLDEN_DIR + '/gae_conc.npz')
    n = len([k for k in z.files if k.endswith('_reward')])
    assert n >= 3
    for i in range(n):
        r, v, d = z['c%d_reward' % i], z['c%d_value' % i], z['c%d_dead' % i]
        ret, threat, pushed = gae(r[:, None, :], v[:, None, :], d[:, None, :], gamma=0.99, tau=0.95, threat_limit=11)
        assert np.array_equal(pushed[:, 0], z['c%d_reward_pushed' % i])
        assert np.array_equal(threat[:, 0], z['c%d_threat' % i])
        assert np.array_equal(ret[:, 0].view(np.uint32), z['c%d_return' % i].view(np.uint32)), i


def test_gae_reward_forwarding_gamma_matches_reference():
    """gamma_in_reward_forwarding (ALGORITHM/conc_4hist/trajectory.py:57-63 = ALGORITHM/common/traj_gae.py:54-60):
    fixtures from the reference's conc_4hist trajectory class with the flag on, forwarding values 0.99 and 0.9."""
    z = np.load(gu.GOLDEN_DIR + '/gae_conc4hist_fwd.npz')
    for tag in 'ab':
        gf = float(z[tag + '_gamma_fwd'])
        n = len([k for k in z.files if k.startswith(tag + '_') and k.endswith('_reward')])
        assert n >= 3
        for i in range(n):
            r, v, d = (z['%s_c%d_%s' % (tag, i, k)] for k in ('reward', 'value', 'dead'))
            ret, threat, pushed = gae(r[:, None, :], v[:, None, :], d[:, None, :], gamma=0.99, tau=0.95, threat_limit=11,
                                      gamma_fwd=gf)
            assert np.array_equal(pushed[:, 0], z['%s_c%d_reward_pushed' % (tag, i)])
            assert np.array_equal(threat[:, 0], z['%s_c%d_threat' % (tag, i)])
            assert np.array_equal(ret[:, 0].view(np.uint32), z['%s_c%d_return' % (tag, i)].view(np.uint32)), (tag, i)


def test_gae_policy_resonance_levels_match_reference():
    """gae_finalize_return_pr (ALGORITHM/common/traj_gae.py:135-169) = the plain scan on G x P columns fed by the pushed
    reward of each agent: fixtures from the reference's common trajectory class with use_policy_resonance / 'level'."""
    z = np.load(gu.GOLDEN_DIR + '/gae_common_pr.npz')
    for i in range(3):
        r, lv, d = z['c%d_reward' % i], z['c%d_levels' % i], z['c%d_dead' % i]
        T, G, P = lv.shape
        _, thr, push = gae(r[:, None], lv[:, None, :, 0], d[:, None], threat_limit=8)
        assert np.array_equal(push[:, 0], z['c%d_reward_pushed' % i]) and np.array_equal(thr[:, 0], z['c%d_threat' % i])
        ret, _, _ = gae(np.repeat(push[:, 0], P, axis=-1)[:, None], lv.reshape(T, 1, G * P), np.zeros((T, 1, G * P), bool),
                        threat_limit=8)
        assert np.array_equal(ret[:, 0].reshape(T, G, P).view(np.uint32), z['c%d_return_levels' % i].view(np.uint32)), i
