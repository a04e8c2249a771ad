"""Diagnostics: per-phase clock breakdown of k_plan (thread-0 clock64 at phase boundaries, averaged over CTAs)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402

NAMES = ['load', 'teams (labels / k-means)', 'guard list + adjacency', 'components', 'targets + means', 'team->target',
         'ring slots', 'member->slot', 'forces']


def main(last=100, window=20, E=4096, preroll=0):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    env = DcaVecEnv(E, scfg=scfg, seed=1)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(0)
    if preroll:  # the stationary episode-phase mix of bench.py: env e sits out the first preroll - offs[e] steps
        offs = torch.randint(0, preroll, (E,), generator=g, device=env.device)
        for s_ in range(preroll):
            act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
            env.step(act, env.plan(), skip=(offs < (preroll - s_)).to(torch.uint8))
    prof = torch.zeros((E, 16), dtype=torch.int64, device=env.device)
    acc = np.zeros(9)
    life, span, tail = [], [], []
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ms = 0.0
    for t in range(last):
        act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
        on = t >= last - window
        env.lib.dca_debug_phase_clocks(env.handle, C.c_void_p(prof.data_ptr()) if on else None)
        ev[0].record()
        f = env.plan()
        ev[1].record()
        env.lib.dca_debug_phase_clocks(env.handle, None)
        env.step(act, f)
        if on:
            torch.cuda.synchronize()
            ms += ev[0].elapsed_time(ev[1])
            c = prof.cpu().numpy().astype(np.float64)
            acc += np.diff(c[:, :10], axis=1).mean(0)
            life.append(c[:, 9] - c[:, 0])
            span.append(c[:, 9].max() - c[:, 0].min())
            if t == last - 1:
                d = np.diff(c[:, :10], axis=1)
                worst = np.argsort(-life[-1])[:5]
                for w in worst:
                    tail.append('env %4d lifetime %7.0f: ' % (w, life[-1][w]) + ' '.join('%6.0f' % x for x in d[w]) +
                                ' | largest team %d, teams > 8: %d (members %d), small teams %d' % tuple(c[w, 10:14]))
                ms_col = d[:, 7]
                for lo_, hi_ in ((0, 8), (9, 12), (13, 16), (17, 22), (23, 32), (33, 999)):
                    sel = (c[:, 10] >= lo_) & (c[:, 10] <= hi_)
                    if sel.any():
                        tail.append('largest team %2d..%3d: %4d envs, member->slot mean %7.0f cycles, lifetime mean %7.0f'
                                    % (lo_, hi_, sel.sum(), ms_col[sel].mean(), life[-1][sel].mean()))
    acc /= window
    for n, v in zip(NAMES, acc):
        print('%-26s %9.0f cycles  %5.1f%%' % (n, v, 100 * v / acc.sum()))
    print('CTA lifetime %.0f cycles, k_plan %.3f ms per launch' % (acc.sum(), ms / window))
    lf = np.concatenate(life)
    print('lifetime percentiles (cycles): p50 %.0f  p90 %.0f  p99 %.0f  max %.0f; first start -> last end per launch %.0f'
          % (np.percentile(lf, 50), np.percentile(lf, 90), np.percentile(lf, 99), lf.max(), np.mean(span)))
    print('slowest envs of the last launch (phase cycles: ' + ', '.join(NAMES) + ')')
    for l in tail:
        print(l)


if __name__ == '__main__':
    main(*(int(a) for a in sys.argv[1:5]))
