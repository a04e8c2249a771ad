"""Throughput of plan_centroid + step issued back to back with and without DCA_OPT_PIPELINE_STEPS (CUDA events)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402


def run(piped, overlap, E=4096, steps=720, planner='centroid'):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    env = DcaVecEnv(E, scfg=scfg, seed=1, overlap_planner=overlap, pipeline_steps=piped)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(0)
    pool = [torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32) for _ in range(16)]
    plan = env.plan_centroid if planner == 'centroid' else env.plan
    for i in range(60):
        env.step(pool[i % 16], plan())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(steps):
        env.step(pool[i % 16], plan())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print('%-9s piped=%d overlap=%d E=%d: %.4f ms per step, %.3f G agent-steps/s' % (planner, piped, overlap, E, ms, E * env.N / ms / 1e6))


if __name__ == '__main__':
    for E in (4096, 8192):
        run(False, False, E); run(False, True, E); run(True, True, E)
    run(False, True, 4096, 360, 'scripted'); run(True, True, 4096, 360, 'scripted')
