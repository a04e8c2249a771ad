"""Diagnostics: wall-clock timeline of the CTAs of consecutive k_step launches in the pipelined loop
(DCA_OPT_PIPELINE_STEPS): per launch first start / last end, CTAs in flight over time summed over the launches, and how
long a CTA sat between its start and the end of its blob load (that span holds the wait for its env's previous step)."""
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
NL = 6


def grouped(groups):
    """The same timeline for `groups` env groups of E / groups envs on their own streams (DcaEnvGroups' arrangement)."""
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    Eg = E // groups
    streams = [torch.cuda.Stream() for _ in range(groups)]
    envs = []
    for k in range(groups):
        with torch.cuda.stream(streams[k]):
            env = DcaVecEnv(Eg, scfg=scfg, seed=1, env_gid0=k * Eg, overlap_planner=True)
            env.reset()
            envs.append(env)
    g = torch.Generator(device=envs[0].device)
    g.manual_seed(0)
    pool = [torch.randint(0, 7, (Eg, envs[0].G), generator=g, device=envs[0].device, dtype=torch.int32) for _ in range(16)]
    profs = [[torch.zeros((Eg, 16), dtype=torch.int64, device=envs[0].device) for _ in range(NL)] for _ in range(groups)]
    torch.cuda.synchronize()
    for t in range(230):
        for k in range(groups):
            with torch.cuda.stream(streams[k]):
                envs[k].step(pool[t % 16], envs[k].plan_centroid())
    for j in range(NL):
        for k in range(groups):
            with torch.cuda.stream(streams[k]):
                ff = envs[k].plan_centroid()
                envs[k].lib.dca_debug_phase_clocks(envs[k].handle, C.c_void_p(profs[k][j].data_ptr()))
                envs[k].step(pool[j % 16], ff)
                envs[k].lib.dca_debug_phase_clocks(envs[k].handle, C.c_void_p(0))
    torch.cuda.synchronize()
    ghz = 1.965
    S, En, which = [], [], []
    for k in range(groups):
        for j in range(NL):
            c = profs[k][j].cpu().numpy().astype(np.uint64)
            life = (c[:, 11] - c[:, 0]).astype(np.float64) / ghz
            end_ns = (c[:, 15] >> np.uint64(8)).astype(np.float64)
            S.append(end_ns - life); En.append(end_ns); which.append(np.full(Eg, k))
    S, En, which = np.concatenate(S), np.concatenate(En), np.concatenate(which)
    t0 = S.min()
    S -= t0; En -= t0
    print('%d groups of %d envs, %d launches each: first start -> last end %.1f us = %.1f us per step of all groups; '
          'CTA lifetime mean %.2f us' % (groups, Eg, NL, En.max() / 1e3, En.max() / 1e3 / NL, (En - S).mean() / 1e3))
    inner = []
    for mid in np.arange(1000.0, En.max(), 4000.0):
        live = (S <= mid) & (En > mid)
        per = ' '.join('%5d' % int((live & (which == k)).sum()) for k in range(groups))
        print('t=%6.1f us  CTAs in flight %5d (%.2f per SM) | per group %s' % (mid / 1e3, int(live.sum()), live.sum() / 148.0, per))
        if 40e3 < mid < En.max() - 40e3:
            inner.append(live.sum())
    inner = np.array(inner)
    print('away from the two ends: mean %.0f of 1184 slots in flight (%.1f %%), minimum %d' % (
        inner.mean(), 100 * inner.mean() / 1184, inner.min()))


if len(sys.argv) > 1 and sys.argv[1] == 'groups':
    grouped(int(sys.argv[2]) if len(sys.argv) > 2 else 2)
    sys.exit(0)

piped = int(sys.argv[1]) if len(sys.argv) > 1 else 1
planner = int(sys.argv[2]) if len(sys.argv) > 2 else 1   # 0: forces fixed (no planner launch between the steps)
scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
env = DcaVecEnv(E, scfg=scfg, seed=1, overlap_planner=True, pipeline_steps=bool(piped))
env.reset()
g = torch.Generator(device=env.device)
g.manual_seed(0)
pool = [torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32) for _ in range(16)]
profs = [torch.zeros((E, 16), dtype=torch.int64, device=env.device) for _ in range(NL)]
f = env.plan_centroid()
for t in range(230):
    env.step(pool[t % 16], env.plan_centroid() if planner else f)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(NL):
    ff = env.plan_centroid() if planner else f
    env.lib.dca_debug_phase_clocks(env.handle, C.c_void_p(profs[k].data_ptr()))
    env.step(pool[k % 16], ff)
    env.lib.dca_debug_phase_clocks(env.handle, C.c_void_p(0))
e1.record()
torch.cuda.synchronize()
print('piped=%d planner=%d: %.1f us per step over %d launches (events)' % (piped, planner, e0.elapsed_time(e1) * 1e3 / NL, NL))
ghz = 1.965
st, en, ld = [], [], []
for k in range(NL):
    c = profs[k].cpu().numpy().astype(np.uint64)
    life = (c[:, 11] - c[:, 0]).astype(np.float64) / ghz
    end_ns = (c[:, 15] >> np.uint64(8)).astype(np.float64)
    st.append(end_ns - life)
    en.append(end_ns)
    ld.append((c[:, 1] - c[:, 0]).astype(np.float64) / ghz)
t0 = min(s.min() for s in st)
for k in range(NL):
    s, e_ = st[k] - t0, en[k] - t0
    print('launch %d: first start %7.1f us  last start %7.1f  first end %7.1f  last end %7.1f | lifetime mean %5.2f us p90 %5.2f max %5.2f | '
          'start->blob loaded mean %5.2f us p90 %5.2f max %5.2f' % (
              k, s.min() / 1e3, s.max() / 1e3, e_.min() / 1e3, e_.max() / 1e3, (e_ - s).mean() / 1e3,
              np.percentile(e_ - s, 90) / 1e3, (e_ - s).max() / 1e3, ld[k].mean() / 1e3, np.percentile(ld[k], 90) / 1e3,
              ld[k].max() / 1e3))
S = np.concatenate(st) - t0
En = np.concatenate(en) - t0
which = np.concatenate([np.full(E, k) for k in range(NL)])
tmax = En.max()
for mid in np.arange(1000.0, tmax, 4000.0):
    live = (S <= mid) & (En > mid)
    per = ' '.join('%5d' % int((live & (which == k)).sum()) for k in range(NL))
    print('t=%6.1f us  CTAs in flight %5d (%.2f per SM) | per launch %s' % (mid / 1e3, int(live.sum()), live.sum() / 148.0, per))
