"""ORACLE / TEST INFRASTRUCTURE -- generates tests/golden/histpool_cases.npz by EXECUTING THE REFERENCE's
ALGORITHM/conc_4hist/shell_env.py ShellEnvWrapper.interact_with_env (:25-66) / solve_duplicate (:68-99) and the
Cython roll_hisory (ALGORITHM/conc_4hist/cython_func.pyx:32-53) on synthetic observation streams.
Run in the build container only (needs /root/reference):

    python oracle/make_golden_histpool.py
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, 'ref_harness'))
sys.path.insert(0, HERE)
OUT = os.path.join(REPO, 'tests', 'golden', 'histpool_cases.npz')


def synth_obs(rng, E, G, dead):
    """DCA-shaped observations: row 0 = self (alive flag 1), other rows present with prob .5, zeros otherwise; a dead
    agent's observation is all zeros (collective_assault_env.py:313-316)."""
    obs = (rng.randint(-7, 8, size=(E, G, 12, 16)) / 4.0).astype(np.float32)  # few distinct values: small fixture
    obs[..., 0] = 1.0
    present = rng.rand(E, G, 12) < 0.5
    present[..., 0] = True
    obs *= present[..., None]
    obs[dead] = 0
    return obs.reshape(E, G, 192)


def run_case(E, G, T, seed, p_done, p_pause):
    from ref_env import setup_reference
    setup_reference()
    from ALGORITHM.conc_4hist.shell_env import ShellEnvWrapper
    rng = np.random.RandomState(seed)
    fed = {}

    class RL(object):
        def interact_with_env_genuine(self, rec):
            fed['obs'] = rec['obs'].copy()
            n = rec['obs'].shape[0]
            return np.zeros((n, G), int), {'_hook_': object()}
    me = types.SimpleNamespace(n_agent=G, n_thread=E, core_dim=16, AvailActProvided=False, RL_functional=RL(),
                               cold_start=True)
    me.solve_duplicate = lambda a, b: ShellEnvWrapper.solve_duplicate(me, a, b)
    state = {'Test-Flag': False, 'Current-Obs-Step': np.zeros(E, int),
             'Latest-Team-Info': np.array([{} for _ in range(E)], dtype=object)}
    dead = np.zeros((E, G), bool)
    rec = {k: [] for k in ('obs', 'pause', 'suffered', 'feed', 'his_pool')}
    for t in range(T):
        P = rng.rand(E) < p_pause if t > 0 else np.zeros(E, bool)
        if P.all():
            P[0] = False
        RST = rng.rand(E) < p_done if t > 0 else np.ones(E, bool)
        dead = (dead | (rng.rand(E, G) < 0.1)) & ~RST[:, None]
        obs = synth_obs(rng, E, G, dead)
        state.update({'Latest-Obs': obs, 'ENV-PAUSE': P, 'Env-Suffered-Reset': RST})
        ShellEnvWrapper.interact_with_env(me, state)
        feed = np.full((E, G, 24, 16), np.nan, np.float32)  # paused envs are not fed: NaN in the uncompacted layout
        feed[~P] = fed['obs']
        rec['obs'].append(obs)
        rec['pause'].append(P)
        rec['suffered'].append(RST)
        rec['feed'].append(feed)
        rec['his_pool'].append(state['_Histpool_Obs_'].copy())
    return {k: np.array(v) for k, v in rec.items()}


if __name__ == '__main__':
    allrec = {}
    for name, args in {'mixed': (5, 7, 30, 1, 0.1, 0.2), 'nopause': (2, 50, 10, 2, 0.1, 0.0)}.items():
        r = run_case(*args)
        for k, v in r.items():
            allrec['%s__%s' % (name, k)] = v
        print(name, {k: v.shape for k, v in r.items()})
    np.savez_compressed(OUT, **allrec)
    print(OUT, '%.1f KB' % (os.path.getsize(OUT) / 1024))
