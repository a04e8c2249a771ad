"""DRAM-counter evidence over a RANGE of back-to-back launches (VERDICT r1 "next" #2).

A single-kernel ncu capture flushes the caches before the launch and stops counting at its end, so obs lines still
dirty in L2 when the kernel ends are never charged and the 28 MB of blobs are charged as cold reads.  Here `--launches`
launches of one kernel run back to back between cudaProfilerStart/Stop and ncu measures the whole range:

    python tools/dram_range.py --kernel observe --launches 64 > plain.log &&
    ncu --replay-mode range --profile-from-start off --cache-control none --clock-control none \
        --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file range_observe.csv \
        python tools/dram_range.py --kernel observe --launches 64

tools/summarize_range.py turns the csv into profiles/latest_ncu.json keys (`k_observe_dram_bytes_range_per_launch` ...),
which bench.py divides by its own CUDA-event duration (`roofline*.dram_frac`).  Without ncu the script prints the
CUDA-event time per launch of the same range.
"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from bench import ENVS_PER_GPU, SEED, WORKLOAD  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402


def main_groups(args):
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    scfg = type('BenchScenarioConfig', (ScenarioConfig,), dict(WORKLOAD))
    n, Eg = args.groups, args.envs // args.groups
    streams = [torch.cuda.Stream(device=dev) for _ in range(n)]
    envs = []
    for k in range(n):
        with torch.cuda.stream(streams[k]):
            env = DcaVecEnv(Eg, scfg=scfg, device=dev, seed=SEED, env_gid0=k * Eg)
            env.lib.dca_set_option(env.handle, 2, 0)
            env.reset()
            envs.append(env)
    lib = envs[0].lib
    sp = [C.c_void_p(s.cuda_stream) for s in streams]
    fp = [C.c_void_p(e._force.data_ptr()) for e in envs]
    pool = [torch.randint(0, 7, (Eg, envs[0].G), device=dev, dtype=torch.int32) for _ in range(8)]
    ap_ = [C.c_void_p(a.data_ptr()) for a in pool]
    torch.cuda.synchronize()
    for i in range(args.warm_steps):
        for k in range(n):
            lib.dca_plan_centroid(envs[k].handle, fp[k], sp[k])
            lib.dca_step(envs[k].handle, ap_[i % 8], fp[k], None, None, 1, sp[k])

    def launch(i):
        for k in range(n):
            if args.kernel == 'observe':
                lib.dca_observe(envs[k].handle, None, sp[k])
            else:
                lib.dca_step(envs[k].handle, ap_[i % 8], fp[k], None, None, 1, sp[k])

    for i in range(5):
        launch(i)
    torch.cuda.synchronize()
    import time
    rt = torch.cuda.cudart()
    t0 = time.perf_counter()
    rt.cudaProfilerStart()
    for i in range(args.launches):
        launch(i)
    rt.cudaProfilerStop()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / args.launches
    print('RANGE kernel=%s groups=%d passes=%d envs=%d ms_per_pass=%.5f (host clock around the synchronised range; not a '
          'bench value under ncu)' % (args.kernel, n, args.launches, n * Eg, ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--kernel', default='observe', choices=['observe', 'step'])
    ap.add_argument('--launches', type=int, default=64)
    ap.add_argument('--envs', type=int, default=ENVS_PER_GPU)
    ap.add_argument('--warm-steps', type=int, default=60, help='env steps before the range (mid-episode state)')
    ap.add_argument('--groups', type=int, default=1,
                    help='the envs as this many independent groups on their own streams (DcaEnvGroups arrangement); '
                         '--launches then counts launches PER GROUP, i.e. passes over all envs')
    args = ap.parse_args()
    if args.groups > 1:
        return main_groups(args)
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    scfg = type('BenchScenarioConfig', (ScenarioConfig,), dict(WORKLOAD))
    env = DcaVecEnv(args.envs, scfg=scfg, device=dev, seed=SEED, env_gid0=0)
    lib, h = env.lib, env.handle
    stream = torch.cuda.current_stream(dev)
    sptr = C.c_void_p(stream.cuda_stream)
    force = torch.zeros((env.E, env.A, 2), dtype=torch.float64, device=dev)
    fptr = C.c_void_p(force.data_ptr())
    pool = [torch.randint(0, 7, (env.E, env.G), device=dev, dtype=torch.int32) for _ in range(8)]
    aptrs = [C.c_void_p(a.data_ptr()) for a in pool]
    lib.dca_set_option(h, 2, 0)  # strictly serial launches: the range holds the one kernel family only
    env.reset()
    for i in range(args.warm_steps):
        lib.dca_plan_centroid(h, fptr, sptr)
        lib.dca_step(h, aptrs[i % 8], fptr, None, None, 1, sptr)
    if args.kernel == 'step':
        lib.dca_plan_centroid(h, fptr, sptr)  # forces stay fixed inside the range (k_plan_centroid is another kernel)

    def launch(i):
        if args.kernel == 'observe':
            return lib.dca_observe(h, None, sptr)
        return lib.dca_step(h, aptrs[i % 8], fptr, None, None, 1, sptr)

    for i in range(5):
        launch(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rt = torch.cuda.cudart()
    e0.record(stream)
    rt.cudaProfilerStart()  # only kernel launches inside the range (range replay admits few other API calls)
    for i in range(args.launches):
        launch(i)
    rt.cudaProfilerStop()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.launches
    print('RANGE kernel=%s launches=%d envs=%d ms_per_launch=%.5f (CUDA events; not a bench value under ncu)'
          % (args.kernel, args.launches, env.E, ms))


if __name__ == '__main__':
    main()
