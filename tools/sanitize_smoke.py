"""Small k_reset / k_step / k_observe workload for compute-sanitizer (memcheck / racecheck / synccheck)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402


def run(G, A, E, steps, max_ep):
    scfg = type('S', (ScenarioConfig,), dict(num_guards=G, num_attackers=A, random_jam_prob=0.05, introduce_terrain=True,
                                            terrain_parameters=[0.05, 0.2], MaxEpisodeStep=max_ep, init_distance=1.0))
    env = DcaVecEnv(E, scfg=scfg, seed=7, debug_outputs=True)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(1)
    for t in range(steps):
        act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
        env.step(act, None)
    env.observe()
    torch.cuda.synchronize()
    print('ok', G, A, E, steps, 'done frac', float(env.done.float().mean()), 'alive', float(env.alive.float().mean()))


if __name__ == '__main__':
    run(50, 50, 48, 30, 12)     # compile-time layout, close formation (laser pairs), several autoresets
    run(100, 100, 16, 12, 6)    # 256-thread kernels
    run(13, 9, 16, 20, 8)       # generic layout, ragged sizes
