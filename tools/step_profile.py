"""Diagnostics: k_step duration as a function of the episode step (bench workload), CUDA events around each launch."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402


def main(steps=380, E=4096):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    env = DcaVecEnv(E, scfg=scfg, seed=1)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(0)
    acts = [torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32) for _ in range(8)]
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    alive = []
    for t in range(steps):
        f = env.plan_centroid()
        ev[t][0].record()
        env.step(acts[t % 8], f)
        ev[t][1].record()
        if t % 10 == 9:
            alive.append(float(env.alive.float().mean().item()))
    torch.cuda.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in ev])
    for k in range(0, steps, 10):
        print('steps %3d-%3d  k_step %.1f us  (min %.1f max %.1f)  alive %.3f' % (k, k + 9, 1e3 * ms[k:k + 10].mean(),
              1e3 * ms[k:k + 10].min(), 1e3 * ms[k:k + 10].max(), alive[k // 10]))
    print('mean over steps 20..379: %.1f us' % (1e3 * ms[20:].mean()))


if __name__ == '__main__':
    main()
