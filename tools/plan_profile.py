"""k_plan launch time over two episodes driven by the scripted planner itself (CUDA events around every launch), with the
slowest envs' phase clocks at the slowest step: where the planner's time goes late in an episode."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402
from plan_phase_clocks import NAMES  # noqa: E402


def main(E=4096, steps=360):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    env = DcaVecEnv(E, scfg=scfg, seed=1, overlap_planner=False)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(0)
    prof = torch.zeros((E, 16), dtype=torch.int64, device=env.device)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    clocks, alive_a, alive_g = [], [], []
    for t in range(steps):
        act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
        env.lib.dca_debug_phase_clocks(env.handle, C.c_void_p(prof.data_ptr()) if t % 10 == 9 else None)
        ev[t][0].record()
        f = env.plan()
        ev[t][1].record()
        env.lib.dca_debug_phase_clocks(env.handle, None)  # k_step writes the same slots
        env.step(act, f)
        if t % 10 == 9:
            torch.cuda.synchronize()
            clocks.append((t, prof.cpu().numpy().astype(np.float64)))
            a = env.alive.float()
            alive_g.append(float(a[:, :env.G].mean())); alive_a.append(float(a[:, env.G:].mean()))
    torch.cuda.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in ev])
    print('k_plan ms per launch: mean %.4f  min %.4f  max %.4f' % (ms.mean(), ms.min(), ms.max()))
    for i, (t, c) in enumerate(clocks):
        life = c[:, 9] - c[:, 0]
        d = np.diff(c[:, :10], axis=1)
        print('step %3d: %.4f ms  alive g %.2f a %.2f  lifetime p50 %6.0f p99 %6.0f max %6.0f | mean phases: %s' % (
            t, ms[t - 9:t + 1].mean(), alive_g[i], alive_a[i], np.percentile(life, 50), np.percentile(life, 99), life.max(),
            ' '.join('%s %.0f' % (n.split(' ')[0], v) for n, v in zip(NAMES, d.mean(0)))))
    t, c = max(clocks, key=lambda tc: (tc[1][:, 9] - tc[1][:, 0]).max())
    life = c[:, 9] - c[:, 0]
    d = np.diff(c[:, :10], axis=1)
    print('slowest envs at step %d (phases: %s)' % (t, ', '.join(NAMES)))
    for w in np.argsort(-life)[:6]:
        print('env %4d lifetime %7.0f: ' % (w, life[w]) + ' '.join('%6.0f' % x for x in d[w]))


if __name__ == '__main__':
    main()
