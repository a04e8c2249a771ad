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


def run(groups, piped, E=4096, steps=720, preroll=True, overlap=True, scripted=False, noplan=False):
    scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
    Eg = E // groups
    streams = [torch.cuda.Stream() for _ in range(groups)]
    envs = []
    for k in range(groups):
        with torch.cuda.stream(streams[k]):
            env = DcaVecEnv(Eg, scfg=scfg, seed=1, env_gid0=k * Eg, overlap_planner=overlap, pipeline_steps=piped)
            env.reset()
            envs.append(env)
    g = torch.Generator(device=envs[0].device)
    g.manual_seed(0)
    pool = [torch.randint(0, 7, (E, envs[0].G), generator=g, device=envs[0].device, dtype=torch.int32) for _ in range(16)]
    pools = [[p[k * Eg:(k + 1) * Eg].contiguous() for p in pool] for k in range(groups)]
    torch.cuda.synchronize()

    import ctypes as C
    lib = envs[0].lib
    sp = [C.c_void_p(s.cuda_stream) for s in streams]
    fp = [C.c_void_p(e._force.data_ptr()) for e in envs]
    ap = [[C.c_void_p(a.data_ptr()) for a in pools[k]] for k in range(groups)]

    def step_all(i):   # raw C-ABI calls, as bench.py's timed loop
        for k in range(groups):
            if noplan:   # forces left as they are: what the step costs with no planner launch in the chain at all
                pass
            elif scripted:
                lib.dca_plan(envs[k].handle, fp[k], None, None, sp[k])
            else:
                lib.dca_plan_centroid(envs[k].handle, fp[k], sp[k])
            lib.dca_step(envs[k].handle, ap[k][i % 16], fp[k], None, None, 1, sp[k])

    if preroll:   # stationary episode-phase mix, as bench.py
        for k in range(groups):
            with torch.cuda.stream(streams[k]):
                gk = torch.Generator(device=envs[k].device)
                gk.manual_seed(100 + k)
                offs = torch.randint(0, 180, (Eg,), generator=gk, device=envs[k].device)
                for s_ in range(180):
                    skip = (offs < (180 - s_)).to(torch.uint8)
                    envs[k].step(pools[k][s_ % 16], envs[k].plan() if scripted else envs[k].plan_centroid(), skip=skip)
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
    print('%s groups=%d piped=%d overlap=%d E=%d: %.4f ms per step of all groups, %.3f G agent-steps/s' % (
        'no planner launch' if noplan else ('scripted' if scripted else 'centroid'), groups, piped, overlap, E, ms,
        E * envs[0].N / ms / 1e6), flush=True)


if __name__ == '__main__':
    if len(sys.argv) > 1 and sys.argv[1] == 'sweep':   # SURVEY.md 8(d) E sweep, one pipelined group against two groups
        for E in (1024, 2048, 4096, 8192, 16384, 32768):
            run(1, True, E=E, steps=360 if E <= 8192 else 180)
            run(2, False, E=E, steps=360 if E <= 8192 else 180)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == 'noplan':
        run(2, False)
        for groups in (2, 3, 4):
            run(groups, False, E=4096 // groups * groups, noplan=True)
        sys.exit(0)
    run(1, True)
    for groups in (2, 4, 8):
        run(groups, False)
        run(groups, False, overlap=False)
    run(2, True)
    run(1, True, steps=360, scripted=True)
    for groups in (2, 4):
        run(groups, True, steps=360, scripted=True)
        run(groups, False, steps=360, scripted=True)
    for E in (2048, 8192):
        run(1, True, E=E)
        run(2, False, E=E)
