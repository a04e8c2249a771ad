"""Diagnostics: per-phase clock breakdown of k_step (thread-0 clock64 at phase boundaries, averaged over CTAs)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402

NAMES = ['load', 'decode+terrain', 'aim', 'laser', 'deaths+integrate', 'reward/done', 'table', 'kNN+shuffle', 'gather',
         'reset path', 'store']


def main(steps=120, window=20, E=4096):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    env = DcaVecEnv(E, scfg=scfg, seed=1)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(0)
    prof = torch.zeros((E, 16), dtype=torch.int64, device=env.device)
    acc = np.zeros(11)
    life = 0.0
    for t in range(steps):
        act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
        on = t >= steps - window
        env.lib.dca_debug_phase_clocks(env.handle, C.c_void_p(prof.data_ptr()) if on else None)
        env.step(act, env.plan_centroid())
        if on:
            c = prof.cpu().numpy().astype(np.float64)
            acc += np.diff(c[:, :12], axis=1).mean(0)
            life += (c[:, 11] - c[:, 0]).mean()
    acc /= window
    life /= window
    for n, v in zip(NAMES, acc):
        print('%-18s %9.0f cycles  %5.1f%%' % (n, v, 100 * v / life))
    print('CTA lifetime       %9.0f cycles' % life)


if __name__ == '__main__':
    # usage: phase_clocks.py [last_step [window]]  -- averages the `window` steps before `last_step` of the episode
    main(*(int(a) for a in sys.argv[1:3])) if len(sys.argv) > 1 else main()
