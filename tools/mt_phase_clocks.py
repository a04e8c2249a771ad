"""Diagnostics: per-phase clock breakdown of k_mt_step (thread-0 clock64 at phase boundaries, averaged over CTAs)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmp2g_b200.multiteam import DcaMultiTeamVecEnv, MtScenarioConfig  # noqa: E402

NAMES = ['load', 'terrain', 'laser', 'deaths+integrate', 'done/ranking', 'obs tasks', 'gather', 'store']


def main(steps=80, E=4096):
    cfg = type('S', (MtScenarioConfig,), dict(N_AGENT_EACH_TEAM=[25, 25], random_jam_prob=0.05, introduce_terrain=True,
                                              terrain_parameters=[0.05, 0.2]))
    env = DcaMultiTeamVecEnv(E, scfg=cfg, seed=1)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(0)
    prof = torch.zeros((E, 16), dtype=torch.int64, device=env.device)
    acc = np.zeros(8)
    life = 0.0
    sub = None
    for t in range(steps):
        act = torch.randint(0, 7, (E, env.N), generator=g, device=env.device, dtype=torch.int32)
        on = t >= steps - 20
        env.lib.dca_mt_debug_phase_clocks(env.handle, C.c_void_p(prof.data_ptr()) if on else None)
        env.step(act)
        if on:
            c = prof.cpu().numpy().astype(np.float64)
            acc += np.diff(c[:, :9], axis=1).mean(0)
            life += (c[:, 8] - c[:, 0]).mean()
            x = np.stack([c[:, 9] - c[:, 2], c[:, 10] - c[:, 9], c[:, 11] - c[:, 10], c[:, 3] - c[:, 11]], 1).mean(0)
            sub = x if sub is None else sub + x
    acc /= 20
    life /= 20
    if sub is not None:
        print('laser sub-phases: bbox %.0f  scan %.0f  barrier %.0f  queue %.0f' % tuple(sub / 20))
    for n, v in zip(NAMES, acc):
        print('%-18s %9.0f cycles  %5.1f%%' % (n, v, 100 * v / life))
    print('CTA lifetime       %9.0f cycles' % life)


if __name__ == '__main__':
    main()
