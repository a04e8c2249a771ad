"""ORACLE / TEST INFRASTRUCTURE -- episode statistics of the UNMODIFIED reference env with its own sklearn / scipy planner
(k-means++ from numpy's MT19937 stream: the one sub-step the restatement reproduces statistically)
under uniform random guard actions: the fixture the keyed planner's statistics are held to
(tests/test_planner.py::test_keyed_planner_episode_statistics_match_reference).  Runs on the staged reference tree
(oracle/stage_ref.py), SmartPool + SuperpoolEnv like the reference's own runner.

    python oracle/make_golden_planner_stats.py 640     # >= 500 episodes -> tests/golden/planner_stats_50v50.json
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import stage_ref  # noqa: E402
from ref_pool_bench import SCENARIOS, inject  # noqa: E402


def main(n_episodes, procs, fold):
    stage_ref.activate()
    import numpy as np
    from config import GlobalConfig as cfg
    E = procs * fold
    cfg.env_name, cfg.num_threads, cfg.fold, cfg.seed = 'dca', E, fold, 31337
    from MISSION.env_router import load_ScenarioConfig, make_parallel_envs
    load_ScenarioConfig()
    inject(cfg.ScenarioConfig, SCENARIOS['50v50'])
    G = cfg.ScenarioConfig.num_guards
    from UTIL.shm_pool import SmartPool
    pool = SmartPool(fold=fold, proc_num=procs, base_seed=31337)
    lens, wins, g_alive = [], [], []
    t0 = time.time()
    try:
        envs = make_parallel_envs(pool)
        envs.reset()
        rng = np.random.RandomState(5)
        cur = np.zeros(E, int)
        while len(lens) < n_episodes:
            ob, rew, done, info = envs.step(rng.randint(0, 7, size=(E, G)).astype(np.float64))
            cur += 1
            for e in np.nonzero(done)[0]:
                lens.append(int(cur[e])); wins.append(bool(info[e]['win']))
                tob = np.asarray(info[e]['obs-echo']).reshape(G, 12, 16)
                g_alive.append(int((tob[:, 0, 0] > 0).sum()))
                cur[e] = 0
    finally:
        pool.party_over()
    lens, wins, g_alive = lens[:n_episodes], wins[:n_episodes], g_alive[:n_episodes]
    out = {'what': 'unmodified reference env + its own planner, uniform random guard actions, 50v50 (bench WORKLOAD), '
                   'first %d finished episodes of %d envs' % (n_episodes, E),
           'episodes': len(lens), 'mean_len': float(np.mean(lens)), 'std_len': float(np.std(lens)),
           'guard_win_rate': float(np.mean(wins)), 'mean_guards_alive_at_end': float(np.mean(g_alive)),
           'std_guards_alive_at_end': float(np.std(g_alive)), 'frac_time_limit': float(np.mean(np.array(lens) >= 180)),
           'seconds': time.time() - t0}
    with open(os.path.join(REPO, 'tests', 'golden', 'planner_stats_50v50.json'), 'w') as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == '__main__':
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 640
    main(n, int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1), int(sys.argv[3]) if len(sys.argv) > 3 else 4)
