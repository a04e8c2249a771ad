"""ORACLE / TEST INFRASTRUCTURE -- generates tests/golden/planner_*.npz by EXECUTING THE REFERENCE planner.

Run in the build container only (needs /root/reference):
    python oracle/make_golden_planner.py

Each file holds, per env step of the unmodified reference env, the inputs and outputs of
CheatScriptAI.attackers_policy_1 (MISSION/dca/cheat_script_ai.py:19-194): guard / attacker positions and alive flags as
the planner saw them, the team labels and k-means centroids it used (the result of get_groups :259-297, i.e. of
scipy.cluster.vq.kmeans2 under numpy's global RNG -- recorded, because that draw sequence is not reproducible
elsewhere), and the resulting attacker.act[0:2].  tests/test_planner.py pins oracle/dca_planner.c to them.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, 'ref_harness'))
sys.path.insert(0, HERE)
OUT = os.path.join(REPO, 'tests', 'golden')

CASES = {
    '50v50': dict(cfg=dict(num_guards=50, num_attackers=50, MaxEpisodeStep=180), steps=110, seed=11, gid=3),
    '50v55': dict(cfg=dict(num_guards=50, num_attackers=55, terrain_parameters=[0.15, 0.2], MaxEpisodeStep=180), steps=60,
                  seed=12, gid=4),
    '20v30_nored': dict(cfg=dict(num_guards=20, num_attackers=30, DISALBE_RED_FUNCTION=True, MaxEpisodeStep=40,
                                 init_distance=2), steps=60, seed=13, gid=5),
}


def run_case(name, spec):
    from ref_env import RefDcaEnv
    env = RefDcaEnv(spec['cfg'], seed=spec['seed'], env_gid=spec['gid'], autoreset=True)
    G, A = env.G, env.A
    world = env.world
    calls = []
    orig = world.get_groups

    def spy(policy_agents, num_members):
        res, cent = orig(policy_agents, num_members)
        calls.append((res, np.array(cent, dtype=np.float64)))
        return res, cent
    world.get_groups = spy
    env.reset()
    rng = np.random.RandomState(spec['seed'] + 1)
    rec = {k: [] for k in ('gpos', 'galive', 'apos', 'aalive', 'label', 'cent', 'force', 'n_calls')}
    skip_next = False
    for t in range(spec['steps']):
        s = env.export_state()
        act = np.full((G,), np.nan) if skip_next else rng.randint(0, 7, size=G).astype(np.float64)
        del calls[:]
        ob, rew, done, info = env.step(act)
        if not skip_next:
            if calls:
                res, cent = calls[-1]
            else:
                res, cent = world.teams_result_step1, np.array(world.team_centroid_step1, dtype=np.float64)
            label = np.full((A,), -1, np.int32)
            for key, agents in res.items():
                for ag in agents:
                    label[ag.iden - G] = int(key)
            assert (label >= 0).all()
            rec['gpos'].append(s['pos'][:G]); rec['galive'].append(s['alive'][:G])
            rec['apos'].append(s['pos'][G:]); rec['aalive'].append(s['alive'][G:])
            rec['label'].append(label); rec['cent'].append(cent)
            # on a done step the env was reset after the planner ran: act of the NEW agents is zero, skip the check there
            rec['force'].append(env.attacker_force() if not done else np.full((A, 2), np.nan))
            rec['n_calls'].append(len(calls))
        skip_next = bool(done) and not skip_next
    out = {k: np.array(v) for k, v in rec.items()}
    out['G'], out['A'] = np.int32(G), np.int32(A)
    out['disable_red'] = np.int32(bool(env.cfg['DISALBE_RED_FUNCTION']))
    path = os.path.join(OUT, 'planner_%s.npz' % name)
    np.savez_compressed(path, **out)
    print(name, 'steps', len(out['label']), 'get_groups calls/step %.2f' % out['n_calls'].mean(),
          'alive guards at end', int(out['galive'][-1].sum()), '%.1f KB' % (os.path.getsize(path) / 1024))


if __name__ == '__main__':
    for n in (sys.argv[1:] or list(CASES)):
        run_case(n, CASES[n])
