"""ORACLE / TEST INFRASTRUCTURE -- times the UNMODIFIED reference CPU path (staged by oracle/stage_ref.py).

SURVEY.md section 8(d) "CPU baseline beside it": SmartPool(proc_num = host cores, fold = E_cpu / proc_num)
(UTIL/shm_pool.pyx:254-424) + SuperpoolEnv (UTIL/shm_env.py:126-168) + EnvAutoReset around the DCA env (MISSION/dca),
made by MISSION.env_router.make_parallel_envs exactly as task_runner.Runner does (task_runner.py:19), random guard
actions float64[E,G] (ALGORITHM/random/foundation.py:28), OMP/MKL/OPENBLAS_NUM_THREADS=1 as main.py:24-29, wall clock
around envs.step only.  The scripted attacker planner runs inside World.step (core.py:189-196), as always in the
reference.  Runs in its own process (SmartPool forks workers; bench.py holds a CUDA context) and prints ONE JSON line.

    python oracle/ref_pool_bench.py --scenario 50v50 --procs 16 --fold 2 --steps 40 --warmup 5
"""
import argparse
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import stage_ref  # noqa: E402

SCENARIOS = {
    '50v50': dict(num_guards=50, num_attackers=50, size=5.0, random_jam_prob=0.05, introduce_terrain=True,
                  terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180, half_death_reward=True, init_distance=4),
    '100v100': dict(num_guards=100, num_attackers=100, size=5.0, random_jam_prob=0.05, introduce_terrain=True,
                    terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180, half_death_reward=True, init_distance=4),
    '50v55': dict(num_guards=50, num_attackers=55, size=5.0, random_jam_prob=0.05, introduce_terrain=True,
                  terrain_parameters=[0.15, 0.2], MaxEpisodeStep=180, half_death_reward=True, init_distance=4,
                  REWARD_DEBUG=True, REWARD_DEBUG_Value=3),
}


def inject(ScenarioConfig, values):
    """What UTIL/config_args.py does for a JSONC group: set the attributes, then recompute the chained twins."""
    for k, v in values.items():
        assert hasattr(ScenarioConfig, k), k
        setattr(ScenarioConfig, k, v)
    G, A = ScenarioConfig.num_guards, ScenarioConfig.num_attackers
    ScenarioConfig.N_AGENT_EACH_TEAM = [G]
    ScenarioConfig.AGENT_ID_EACH_TEAM = [range(0, G)]
    ScenarioConfig.num_entity = A
    ScenarioConfig.num_object = G + A
    ScenarioConfig.uid_dictionary = {'agent_uid': range(0, G), 'entity_uid': range(G, G + A),
                                     'guards_uid': range(0, G), 'attackers_uid': range(G, G + A)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--scenario', default='50v50', choices=sorted(SCENARIOS))
    ap.add_argument('--procs', type=int, default=os.cpu_count() or 1)
    ap.add_argument('--fold', type=int, default=2)
    ap.add_argument('--steps', type=int, default=40)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--seed', type=int, default=20240)
    args = ap.parse_args()
    stage_ref.activate()
    import numpy as np
    from config import GlobalConfig as cfg
    E = args.procs * args.fold
    cfg.env_name = 'dca'
    cfg.num_threads = E
    cfg.fold = args.fold
    cfg.seed = args.seed
    # config injection happens before the worker processes are forked, as in main.py:68-79 (prepare_args, then SmartPool)
    from MISSION.env_router import load_ScenarioConfig, make_parallel_envs
    load_ScenarioConfig()
    inject(cfg.ScenarioConfig, SCENARIOS[args.scenario])
    G, N = cfg.ScenarioConfig.num_guards, cfg.ScenarioConfig.num_object
    from UTIL.shm_pool import SmartPool
    pool = SmartPool(fold=args.fold, proc_num=args.procs, base_seed=args.seed)
    out = None
    try:
        envs = make_parallel_envs(pool)
        envs.reset()
        rng = np.random.RandomState(1)
        acts = [rng.randint(0, 7, size=(E, G)).astype(np.float64) for _ in range(8)]
        for i in range(args.warmup):
            envs.step(acts[i % 8])
        n_done = 0
        t0 = time.perf_counter()
        for i in range(args.steps):
            ob, rew, done, info = envs.step(acts[i % 8])
            n_done += int(done.sum())
        dt = time.perf_counter() - t0
        out = {'impl': 'reference', 'path': 'SmartPool + SuperpoolEnv + EnvAutoReset + MISSION/dca (unmodified, staged)',
               'scenario': args.scenario, 'procs': args.procs, 'fold': args.fold, 'envs': E, 'steps': args.steps,
               'warmup': args.warmup, 'seconds': dt, 'env_steps_per_sec': E * args.steps / dt,
               'agent_steps_per_sec': E * N * args.steps / dt, 'ms_per_step': dt / args.steps * 1e3,
               'obs_shape': list(ob.shape), 'obs_dtype': str(ob.dtype), 'episodes_ended': n_done,
               'host_cores': os.cpu_count()}
    finally:
        pool.party_over()
    sys.stdout.flush()
    print('REF_POOL_JSON ' + json.dumps(out), flush=True)


if __name__ == '__main__':
    main()
