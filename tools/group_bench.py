"""4096 envs per GPU as 1 / 2 / 4 independent env groups on their own streams (each a DcaVecEnv of E / groups envs, global env
ids contiguous): the launch chains of the groups are independent, so one group's CTAs fill the slots another group's
launch boundary leaves empty (tools/pipe_timeline.py: 4.8 us of slot time per step with one group)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402


def run(groups, piped, E=4096, steps=720, preroll=True):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    Eg = E // groups
    streams = [torch.cuda.Stream() for _ in range(groups)]
    envs = []
    for k in range(groups):
        with torch.cuda.stream(streams[k]):
            env = DcaVecEnv(Eg, scfg=scfg, seed=1, env_gid0=k * Eg, overlap_planner=True, pipeline_steps=piped)
            env.reset()
            envs.append(env)
    g = torch.Generator(device=envs[0].device)
    g.manual_seed(0)
    pool = [torch.randint(0, 7, (E, envs[0].G), generator=g, device=envs[0].device, dtype=torch.int32) for _ in range(16)]
    pools = [[p[k * Eg:(k + 1) * Eg].contiguous() for p in pool] for k in range(groups)]
    torch.cuda.synchronize()

    def step_all(i):
        for k in range(groups):
            with torch.cuda.stream(streams[k]):
                envs[k].step(pools[k][i % 16], envs[k].plan_centroid())

    if preroll:   # stationary episode-phase mix, as bench.py
        for k in range(groups):
            with torch.cuda.stream(streams[k]):
                gk = torch.Generator(device=envs[k].device)
                gk.manual_seed(100 + k)
                offs = torch.randint(0, 180, (Eg,), generator=gk, device=envs[k].device)
                for s_ in range(180):
                    skip = (offs < (180 - s_)).to(torch.uint8)
                    envs[k].step(pools[k][s_ % 16], envs[k].plan_centroid(), skip=skip)
    for i in range(200):
        step_all(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(groups):
        streams[k].wait_event(e0)
    for i in range(steps):
        step_all(i)
    ends = []
    for k in range(groups):
        ev = torch.cuda.Event()
        ev.record(streams[k])
        ends.append(ev)
    for ev in ends:
        torch.cuda.current_stream().wait_event(ev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print('groups=%d piped=%d E=%d: %.4f ms per step of all groups, %.3f G agent-steps/s' % (
        groups, piped, E, ms, E * envs[0].N / ms / 1e6), flush=True)


if __name__ == '__main__':
    for groups in (1, 2, 4):
        for piped in (True, False):
            run(groups, piped)
