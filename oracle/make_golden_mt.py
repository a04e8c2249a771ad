"""ORACLE / TEST INFRASTRUCTURE -- generates tests/golden/mt_*.npz by EXECUTING THE REFERENCE's dca_multiteam env
(MISSION/dca_multiteam via UTIL/shm_env.EnvAutoReset) under the keyed RNG stream.  Run in the build container only:

    python oracle/make_golden_mt.py
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
    # the env's defaults (25 v 25, no terrain, no jamming) with a shorter episode: time-limit done + autoreset inside
    '2x25': dict(cfg=dict(N_AGENT_EACH_TEAM=[25, 25], MaxEpisodeStep=60, StepsToNarrowCircle=40), steps=80, seed=11, gid=4),
    # terrain advantage in the hit probability, jamming, three uneven teams, fast kill circle, full death penalty
    '3teams': dict(cfg=dict(N_AGENT_EACH_TEAM=[7, 9, 8], random_jam_prob=0.2, introduce_terrain=True,
                            terrain_parameters=[0.05, 0.2], MaxEpisodeStep=45, StepsToNarrowCircle=15, size=3.0,
                            init_distance=1.0, half_death_reward=False), steps=110, seed=7, gid=123456),
    # four small teams (6 is the smallest team the reference's observation supports) starting close: early annihilations (teams_alive <= 1 endings), draws in the ranking
    '4x6': dict(cfg=dict(N_AGENT_EACH_TEAM=[6, 6, 6, 6], random_jam_prob=0.05, introduce_terrain=True,
                         terrain_parameters=[0.15, 0.2], MaxEpisodeStep=30, StepsToNarrowCircle=10, size=2.0,
                         init_distance=0.6), steps=100, seed=3, gid=9),
    # the same four teams with the team-level interface: one reward per team (0, then the ranking bonus on done) and the
    # observation of agent 0 only (collective_assault_parallel_run.py:166-186)
    '4x6_unity': dict(cfg=dict(N_AGENT_EACH_TEAM=[6, 6, 6, 6], random_jam_prob=0.05, introduce_terrain=True,
                               terrain_parameters=[0.15, 0.2], MaxEpisodeStep=30, StepsToNarrowCircle=10, size=2.0,
                               init_distance=0.6, RewardAsUnity=True, ObsAsUnity=True), steps=100, seed=3, gid=9),
}
STATE_KEYS = ('pos', 'vel', 'atk_rad', 'terrain', 'alive', 'just_died', 'hit', 'was_hit', 'bsa', 'init_theta',
              'time_step', 'obs_counter', 'episode')


def run_case(name, spec):
    from ref_mt_env import RefMtEnv
    env = RefMtEnv(spec['cfg'], seed=spec['seed'], env_gid=spec['gid'])
    N, T = env.N, len(env.cfg['N_AGENT_EACH_TEAM'])
    obs0, _ = env.reset()
    rec = {'cfg_json': json.dumps(env.cfg), 'seed': spec['seed'], 'gid': spec['gid'], 'obs0': obs0.astype(np.float32)}
    s0 = env.export_state()
    for k in STATE_KEYS:
        rec['s0_' + k] = s0[k]
    rng = np.random.RandomState(spec['seed'] + 1)
    per = {k: [] for k in STATE_KEYS}
    per.update(actions=[], obs=[], reward=[], done=[], ranking=[])
    tobs_steps, tobs = [], []
    skip_next = False
    for t in range(spec['steps']):
        act = np.full((N,), np.nan) if skip_next else rng.randint(0, 7, size=N).astype(np.float64)
        ob, rew, done, info = env.step(act)
        was_skip = skip_next
        skip_next = bool(done) and not skip_next
        per['actions'].append(act)
        per['obs'].append(ob.astype(np.float32))
        assert np.array_equal(ob.astype(np.float32).astype(ob.dtype), ob)
        per['reward'].append(rew)
        per['done'].append(done)
        per['ranking'].append(np.asarray(info['team_ranking'], np.int32) if done else np.full((T,), -9, np.int32))
        if done and not was_skip:
            assert set(info) == {'team_ranking', 'end_reason', 'obs-echo'}, set(info)
            tobs_steps.append(t)
            tobs.append(np.asarray(info['obs-echo'], dtype=np.float32))
        s = env.export_state()
        for k in STATE_KEYS:
            per[k].append(s[k])
    for k, v in per.items():
        rec[k] = np.array(v)
    rec['done'] = rec['done'].astype(np.uint8)
    rec['tobs_steps'] = np.array(tobs_steps, np.int32)
    rec['tobs'] = np.array(tobs, np.float32).reshape(len(tobs), -1, 204)
    path = os.path.join(OUT, 'mt_%s.npz' % name)
    np.savez_compressed(path, **rec)
    print(name, 'steps', spec['steps'], 'dones at', np.nonzero(rec['done'])[0].tolist(), 'rankings',
          [rec['ranking'][t].tolist() for t in tobs_steps], 'alive end', int(rec['alive'][-1].sum()),
          'min alive', int(rec['alive'].sum(1).min()), '%.1f KB' % (os.path.getsize(path) / 1024))


if __name__ == '__main__':
    for n in (sys.argv[1:] or list(CASES)):
        run_case(n, CASES[n])
