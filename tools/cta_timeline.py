"""Diagnostics: wall-clock timeline of the CTAs of one k_step launch (globaltimer at CTA end + clock64 lifetime, SM id):
how many CTAs are in flight over the launch, per-SM busy time, gaps."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402

E = 4096
scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
env = DcaVecEnv(E, scfg=scfg, seed=1, overlap_planner=False)
env.reset()
g = torch.Generator(device=env.device)
g.manual_seed(0)
prof = torch.zeros((E, 16), dtype=torch.int64, device=env.device)
for t in range(int(sys.argv[1]) if len(sys.argv) > 1 else 110):
    act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
    f = env.plan_centroid()
    env.lib.dca_debug_phase_clocks(env.handle, C.c_void_p(prof.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    env.step(act, f)
    e1.record()
torch.cuda.synchronize()
c = prof.cpu().numpy().astype(np.uint64)
life_cyc = (c[:, 11] - c[:, 0]).astype(np.float64)
end_ns = (c[:, 15] >> np.uint64(8)).astype(np.float64)
sm = (c[:, 15] & np.uint64(0xff)).astype(np.int64)
ghz = 1.965
start_ns = end_ns - life_cyc / ghz
t0 = start_ns.min()
start_ns -= t0
end_ns -= t0
print('event time of the launch %.1f us; first CTA start -> last CTA end %.1f us' % (e0.elapsed_time(e1) * 1e3, end_ns.max() / 1e3))
print('CTA lifetime: mean %.2f us  p10 %.2f  p50 %.2f  p90 %.2f  max %.2f' % tuple(
    x / 1e3 for x in (life_cyc.mean() / ghz, np.percentile(life_cyc, 10) / ghz, np.percentile(life_cyc, 50) / ghz,
                      np.percentile(life_cyc, 90) / ghz, life_cyc.max() / ghz)))
print('SMs seen %d; CTAs per SM min %d max %d' % (len(np.unique(sm)), np.bincount(sm).min(), np.bincount(sm).max()))
edges = np.linspace(0, end_ns.max(), 41)
for a, b in zip(edges[:-1], edges[1:]):
    mid = 0.5 * (a + b)
    n = int(((start_ns <= mid) & (end_ns > mid)).sum())
    print('t=%6.1f us  CTAs in flight %5d  (%.2f per SM)' % (mid / 1e3, n, n / 148.0))
# order of launch: start time by block index
order = np.argsort(start_ns)
print('block index of the first 8 / last 8 CTAs to start:', order[:8], order[-8:])
busy = np.zeros(148)
for s_ in range(148):
    sel = sm == s_
    if sel.any():
        busy[s_] = end_ns[sel].max() - start_ns[sel].min()
print('per-SM span (first start -> last end): min %.1f us  mean %.1f  max %.1f' % (busy.min() / 1e3, busy.mean() / 1e3, busy.max() / 1e3))
gaps = []
for s_ in range(148):
    sel = np.nonzero(sm == s_)[0]
    st = np.sort(start_ns[sel]); en = np.sort(end_ns[sel])
    # time between a CTA ending and the next one starting on the same SM (k-th end vs (k+8)-th start)
    if len(st) > 8:
        gaps.extend(list(st[8:] - en[:len(st) - 8]))
gaps = np.array(gaps)
print('end -> replacement start on the same SM: mean %.2f us  p50 %.2f  p90 %.2f' % (gaps.mean() / 1e3, np.percentile(gaps, 50) / 1e3, np.percentile(gaps, 90) / 1e3))
NAMES = ['load', 'decode+terrain', 'aim', 'laser', 'deaths+integrate', 'reward/done', 'table', 'kNN+shuffle', 'gather',
         'reset path', 'store']
done = env.done.cpu().numpy()
top = np.argsort(-life_cyc)[:6]
ph = np.diff(c[:, :12].astype(np.float64), axis=1)
print('phase cycles of the 6 longest CTAs (and the mean over all):')
print('%-18s' % 'mean', ' '.join('%7.0f' % v for v in ph.mean(0)))
for e_ in top:
    print('env %4d done=%d    ' % (e_, done[e_]), ' '.join('%7.0f' % v for v in ph[e_]), ' start %.1f us' % (start_ns[e_] / 1e3),
          ' sincos %d cyc (atk hi word %08x)  thread0 C1 %d cyc  barrier wait %d cyc' % (c[e_, 12] & 0xffffffff, c[e_, 12] >> 32, c[e_, 13], c[e_, 14]))
#print('all envs: shooters mean %.1f max %d; queued pairs mean %.1f max %d' % (c[:, 12].mean(), c[:, 12].max(), c[:, 13].mean(), c[:, 13].max()))
print('done envs this step: %d; their mean lifetime %.1f us; others %.1f us' % (
    done.sum(), life_cyc[done != 0].mean() / ghz / 1e3 if done.any() else 0, life_cyc[done == 0].mean() / ghz / 1e3))
