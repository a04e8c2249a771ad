"""ORACLE / TEST INFRASTRUCTURE -- generates tests/golden/*.npz by EXECUTING THE REFERENCE.

Run in the build container only (needs /root/reference):
    python oracle/make_golden.py            # writes tests/golden/dca_<name>.npz + gae_conc.npz + gae_conc4hist_fwd.npz

Each DCA file holds one env trajectory of the unmodified reference env
(MISSION/dca via UTIL/shm_env.EnvAutoReset) under the keyed RNG stream: initial state after reset,
per-step guard actions, the planner's attacker forces (recorded, cheat_script_ai.py:183-184),
and per-step outputs + full entity state.  tests/test_oracle_golden.py pins oracle/dca_oracle.c to them;
the GPU parity tests replay the same inputs through the C-ABI.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, 'ref_harness'))
sys.path.insert(0, HERE)
OUT = os.path.join(REPO, 'tests', 'golden')

CASES = {
    # BASELINE configs 1/2 (RESULT/test-50+50/test-50+50.jsonc:23-35), shortened episode so a time-limit
    # done + autoreset falls inside the 100-step horizon
    'cfgA_50v50': dict(cfg=dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True,
                                terrain_parameters=[0.05, 0.2], MaxEpisodeStep=80), steps=100, seed=20231, gid=5),
    # BASELINE config 4 (RESULT/50RL-55opp/test-50RL-55opp.jsonc:23-36)
    'cfgC_50v55': dict(cfg=dict(num_guards=50, num_attackers=55, random_jam_prob=0.05, introduce_terrain=True,
                                terrain_parameters=[0.15, 0.2], REWARD_DEBUG=True, REWARD_DEBUG_Value=3,
                                MaxEpisodeStep=180), steps=100, seed=22333, gid=0),
    # BASELINE config 3 (RESULT/test-100+100/test-100+100.jsonc:26-27)
    'cfgB_100v100': dict(cfg=dict(num_guards=100, num_attackers=100, random_jam_prob=0.05, introduce_terrain=True,
                                  terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180), steps=40, seed=3, gid=17),
    # edge cases: attacker fluctuation, tight walls (soft wall force active), heavy jamming, full death penalty,
    # short episodes (several autoresets, all-dead terminations)
    'small_8v10_fluct': dict(cfg=dict(num_guards=8, num_attackers=10, random_jam_prob=0.3, introduce_terrain=False,
                                      enable_attacker_fluctuation=True, attacker_fluctuation=3, size=2.0,
                                      half_death_reward=False, MaxEpisodeStep=40, init_distance=2), steps=130,
                             seed=99, gid=1234567),
    'small_6v10_nored': dict(cfg=dict(num_guards=6, num_attackers=10, random_jam_prob=0.0, introduce_terrain=True,
                                     terrain_parameters=[0.05, 0.2], DISALBE_RED_FUNCTION=True, MaxEpisodeStep=30,
                                     init_distance=1), steps=70, seed=1, gid=2),
}

STATE_KEYS = ('pos', 'vel', 'atk_rad', 'terrain', 'alive', 'just_died', 'hit', 'was_hit', 'num_hit', 'num_was_hit',
              'init_theta', 'time_step', 'n_alive_guards', 'n_alive_attackers', 'obs_counter', 'episode')


def nearest_from_reference(env):
    """argmin row of the cached distance matrix exactly as cheat_script_ai.py:200-207 reads it."""
    w = env.world
    dis = w.shared_resorce['distance_matrix']
    G, A = env.G, env.A
    out = np.full((A,), -1, np.int32)
    for a in range(A):
        if w.agents[G + a].alive:
            out[a] = int(np.argmin(dis[G + a, :G]))
    return out


def run_case(name, spec):
    from ref_env import RefDcaEnv
    env = RefDcaEnv(spec['cfg'], seed=spec['seed'], env_gid=spec['gid'], autoreset=True)
    G, A = env.G, env.A
    obs0, _ = env.reset()
    rec = {'cfg_json': json.dumps(env.cfg), 'seed': spec['seed'], 'gid': spec['gid'],
           'obs0': obs0.astype(np.float32)}
    s0 = env.export_state()
    for k in STATE_KEYS:
        rec['s0_' + k] = s0[k]
    act_rng = np.random.RandomState(spec['seed'] + 1)
    per = {k: [] for k in STATE_KEYS}
    per.update(actions=[], att_force=[], obs=[], reward=[], done=[], win=[], nearest=[])
    tobs_steps, tobs = [], []
    skip_next = False
    prev = None
    for t in range(spec['steps']):
        if skip_next:
            act = np.full((G,), np.nan)  # ENV-PAUSE row (multi_team.py:62-66)
        else:
            act = act_rng.randint(0, 7, size=G).astype(np.float64)
        near = nearest_from_reference(env)
        ob, rew, done, info = env.step(act)
        if skip_next:
            assert prev is not None and np.array_equal(ob, prev[0]) and done == prev[2]
            near[:] = -2  # not evaluated on a skipped step
        skip_next = bool(done) and not skip_next
        prev = (ob, rew, done)
        per['actions'].append(act)
        per['att_force'].append(env.attacker_force())
        per['obs'].append(ob.astype(np.float32))
        assert np.array_equal(ob.astype(np.float32).astype(ob.dtype), ob)
        per['reward'].append(rew)
        per['done'].append(done)
        per['win'].append(bool(info.get('win', False)))
        per['nearest'].append(near)
        if done and 'obs-echo' in info and not np.isnan(act).any():
            tobs_steps.append(t)
            tobs.append(np.asarray(info['obs-echo'], dtype=np.float32))
        s = env.export_state()
        for k in STATE_KEYS:
            per[k].append(s[k])
    for k, v in per.items():
        rec[k] = np.array(v)
    rec['done'] = rec['done'].astype(np.uint8)
    rec['win'] = rec['win'].astype(np.uint8)
    rec['tobs_steps'] = np.array(tobs_steps, np.int32)
    rec['tobs'] = np.array(tobs, np.float32).reshape(len(tobs), G, 192)
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, 'dca_%s.npz' % name)
    np.savez_compressed(path, **rec)
    print(name, 'steps', spec['steps'], 'dones at', np.nonzero(rec['done'])[0].tolist(),
          'alive end', int(rec['alive'][-1].sum()), '%.1f KB' % (os.path.getsize(path) / 1024))


def _load_trajectory_module(alg, **cfg):
    """Import ALGORITHM/<alg>/trajectory.py alone: its `.foundation` (torch-heavy) is replaced by a stub module that
    carries the AlgorithmConfig numbers trajectory.py reads (ALGORITHM/conc/foundation.py:9-10,
    ALGORITHM/conc_4hist/foundation.py:15-16,45-46)."""
    from ref_env import setup_reference
    setup_reference()
    import importlib
    import types
    stub = types.ModuleType('ALGORITHM.%s.foundation' % alg)

    class AlgorithmConfig:
        gamma = 0.99
        tau = 0.95
    for k, v in cfg.items():
        setattr(AlgorithmConfig, k, v)
    stub.AlgorithmConfig = AlgorithmConfig
    pkg = types.ModuleType('ALGORITHM.%s' % alg)
    pkg.__path__ = [os.path.join('/root/reference', 'ALGORITHM', alg)]
    sys.modules['ALGORITHM.%s' % alg] = pkg
    sys.modules['ALGORITHM.%s.foundation' % alg] = stub
    sys.modules.pop('ALGORITHM.%s.trajectory' % alg, None)  # re-import under the new AlgorithmConfig
    return importlib.import_module('ALGORITHM.%s.trajectory' % alg), AlgorithmConfig


def _gae_cases(traj_mod, shapes, seed, rewards=(0.0, 1.0, -0.5, 0.5)):
    rng = np.random.RandomState(seed)
    recs = {}
    for idx, (T, G) in enumerate(shapes):
        tr = traj_mod.trajectory(traj_limit=T, env_id=0)
        # dead mask: agents die at random times and stay dead (obs rows all-NaN once dead)
        death = rng.randint(1, T + 30, size=G)
        obs = rng.randn(T, G, 12, 16).astype(np.float32)
        for g in range(G):
            obs[death[g]:, g] = np.nan
        reward = rng.choice(list(rewards), size=(T, G)).astype(np.float64)
        value = rng.randn(T, G, 1).astype(np.float32)
        tr.obs, tr.reward, tr.value = obs, reward.copy(), value
        tr.key_dict = ['obs', 'reward', 'value']
        tr.time_pointer = T
        tr.finalize()
        recs['c%d_reward' % idx] = reward
        recs['c%d_value' % idx] = value[..., 0]
        recs['c%d_dead' % idx] = np.isnan(obs.reshape(T, G, -1)).all(-1)
        recs['c%d_return' % idx] = getattr(tr, 'return')[..., 0]
        recs['c%d_threat' % idx] = tr.threat[..., 0].astype(np.float32)
        recs['c%d_reward_pushed' % idx] = tr.reward
    return recs


def run_gae():
    """GAE fixtures from ALGORITHM/conc/trajectory.py (conc) run on synthetic trajectories."""
    traj_mod, _ = _load_trajectory_module('conc')
    recs = _gae_cases(traj_mod, [(57, 50), (180, 50), (1, 6), (23, 100)], seed=7)
    path = os.path.join(OUT, 'gae_conc.npz')
    np.savez_compressed(path, **recs)
    print('gae_conc', '%.1f KB' % (os.path.getsize(path) / 1024))


def run_gae_fwd():
    """gamma_in_reward_forwarding fixtures from ALGORITHM/conc_4hist/trajectory.py:57-63 (SAFE_LIMIT 11): a dead
    frame's reward moves to the previous frame scaled by gamma_in_reward_forwarding_value."""
    recs = {}
    for tag, gfv in (('a', 0.99), ('b', 0.9)):
        traj_mod, cfg = _load_trajectory_module('conc_4hist', gamma_in_reward_forwarding=True,
                                                gamma_in_reward_forwarding_value=gfv)
        assert cfg.gamma_in_reward_forwarding and traj_mod.AlgorithmConfig is cfg
        r = _gae_cases(traj_mod, [(57, 50), (180, 20), (2, 6)], seed=11 if tag == 'a' else 12,
                       rewards=(0.0, 1.0, -0.5, 0.5, 0.3, -1.7))
        for k, v in r.items():
            recs[tag + '_' + k] = v
        recs[tag + '_gamma_fwd'] = np.float64(gfv)
    path = os.path.join(OUT, 'gae_conc4hist_fwd.npz')
    np.savez_compressed(path, **recs)
    print('gae_conc4hist_fwd', '%.1f KB' % (os.path.getsize(path) / 1024))


def run_gae_pr():
    """Policy-resonance ('level') fixtures from ALGORITHM/common/traj_gae.py: finalize() with use_policy_resonance and
    policy_resonance_method == 'level' (:86-90) -> gae_finalize_return_pr (:135-169): the GAE scan on every one of the
    `distribution_precision` value levels of an agent, all fed by the agent's pushed-forward reward; SAFE_LIMIT 8."""
    from ref_env import setup_reference
    setup_reference()
    import importlib
    mod = importlib.import_module('ALGORITHM.common.traj_gae')

    class Cfg:
        gamma, tau = 0.99, 0.95
        gamma_in_reward_forwarding, gamma_in_reward_forwarding_value = False, 0.99
        use_policy_resonance, policy_resonance_method, distribution_precision = True, 'level', 5
    rng = np.random.RandomState(21)
    recs = {}
    for idx, (T, G) in enumerate([(40, 6), (120, 20), (3, 7)]):
        mod.trajectory.dead_mask_check = False
        tr = mod.trajectory(traj_limit=T, env_id=0, alg_cfg=Cfg)
        death = rng.randint(1, T + 15, size=G)
        alive = np.arange(T)[:, None] < death[None]
        reward = rng.choice([0.0, 1.0, -0.5, 0.5, 2.0], size=(T, G)).astype(np.float64)
        levels = rng.randn(T, G, Cfg.distribution_precision).astype(np.float32)
        tr.alive, tr.reward, tr.BAL_value_all_level = alive, reward.copy(), levels
        tr.randl = rng.randint(0, Cfg.distribution_precision, size=T)
        tr.key_dict = ['alive', 'reward', 'BAL_value_all_level']
        tr.time_pointer = T
        tr.finalize()
        recs['c%d_reward' % idx] = reward
        recs['c%d_levels' % idx] = levels
        recs['c%d_dead' % idx] = ~alive
        recs['c%d_return_levels' % idx] = tr.BAL_return_all_level
        recs['c%d_threat' % idx] = tr.threat[..., 0].astype(np.float32)
        recs['c%d_reward_pushed' % idx] = tr.reward
    path = os.path.join(OUT, 'gae_common_pr.npz')
    np.savez_compressed(path, **recs)
    print('gae_common_pr', '%.1f KB' % (os.path.getsize(path) / 1024))


if __name__ == '__main__':
    which = sys.argv[1:] or (list(CASES) + ['gae', 'gae_fwd', 'gae_pr'])
    for n in which:
        if n == 'gae':
            run_gae()
        elif n == 'gae_fwd':
            run_gae_fwd()
        elif n == 'gae_pr':
            run_gae_pr()
        else:
            run_case(n, CASES[n])
