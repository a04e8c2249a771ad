"""ORACLE / TEST INFRASTRUCTURE -- statistical parity of the keyed planner (run in the build container only).

Episodes of the unmodified reference env (its own sklearn/scipy planner under numpy's RNG) against episodes of the C
oracle driven by the keyed planner restatement (oracle/dca_planner.c), both with uniform random guard actions:
episode length, attacker win rate and survivors at the end.  Prints one JSON line (copied into DESIGN.md).
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'ref_harness'))
sys.path.insert(0, HERE)

CFG = dict(num_guards=50, num_attackers=50, random_jam_prob=0.05, introduce_terrain=True, terrain_parameters=[0.05, 0.2],
           MaxEpisodeStep=180)


def reference_episodes(n_ep, seed):
    from ref_env import RefDcaEnv
    env = RefDcaEnv(CFG, seed=seed, env_gid=0, autoreset=False, np_seed=seed)
    rng = np.random.RandomState(seed)
    out = []
    for ep in range(n_ep):
        env.reset()
        for t in range(CFG['MaxEpisodeStep']):
            ob, rew, done, info = env.step(rng.randint(0, 7, size=env.G).astype(np.float64))
            if done:
                s = env.export_state()
                out.append((t + 1, bool(info.get('win', False)), int(s['alive'][:env.G].sum()), int(s['alive'][env.G:].sum())))
                break
    return out


def oracle_episodes(n_env, steps, seed):
    from dca_oracle import OraclePlanner, OracleVecEnv
    o = OracleVecEnv(CFG, n_env, seed=seed, env_gid0=0, nthreads=os.cpu_count())
    o.reset()
    pl = OraclePlanner(o)
    rng = np.random.RandomState(seed)
    out = []
    length = np.zeros((n_env,), np.int64)
    for t in range(steps):
        pre_alive = o.state['alive'].copy()
        o.step(rng.randint(0, 7, size=(n_env, o.G)).astype(np.float64), pl.plan(), autoreset=True)
        length += 1
        for e in np.nonzero(o.done)[0]:
            # survivors at the terminal step: alive before the step minus this step's deaths are not exported after an
            # autoreset, so count from the terminal obs' alive column instead
            tob = o.tobs[e].reshape(o.G, 12, 16)
            out.append((int(length[e]), bool(o.win[e]), int((tob[:, 0, 0] > 0).sum()), -1))
            length[e] = 0
    return out


def summarize(eps):
    a = np.array([(l, w, g) for l, w, g, _ in eps], dtype=np.float64)
    return dict(episodes=len(eps), mean_len=a[:, 0].mean(), std_len=a[:, 0].std(), guard_win_rate=a[:, 1].mean(),
                mean_guards_alive_at_end=a[:, 2].mean())


if __name__ == '__main__':
    n_ref = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    t0 = time.time()
    ref = summarize(reference_episodes(n_ref, 1))
    t1 = time.time()
    ours = summarize(oracle_episodes(256, 400, 1))
    print(json.dumps(dict(reference=ref, keyed_planner=ours, ref_seconds=t1 - t0, ours_seconds=time.time() - t1)))
