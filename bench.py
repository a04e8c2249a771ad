"""bench.py -- DCA 50v50 step+obs throughput (agent-steps/s) on N B200s, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: the planner stand-in kernel (attacker forces) + the fused
step/observation/autoreset kernel for every env of this rank.  Workload (BASELINE.json configs[1]): DCA 50v50, 4096
parallel envs per GPU, random guard actions injected from a pre-generated device pool, terrain on, jam 0.05.
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for every definition.
"""
import argparse
import json
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

import numpy as np  # noqa: E402

WORKLOAD = dict(num_guards=50, num_attackers=50, size=5.0, random_jam_prob=0.05, introduce_terrain=True,
                terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180, half_death_reward=True, init_distance=4)
ENVS_PER_GPU = 4096
SEED = 20240
# algorithmic bytes per env-step, 50v50 = obs 38400 + state r/w 12000 + actions 200 +
# attacker force 800 + reward 200 + done/win/alive 102
BYTES_PER_ENV_STEP = 38400 + 12000 + 200 + 800 + 200 + 102
SCENARIOS = {  # BASELINE.json configs: [1] 50v50 (default, the headline), [2] 100v100, [3] 50 RL vs 55 scripted
    '50v50': {},
    '100v100': dict(num_guards=100, num_attackers=100),
    '50v55': dict(num_guards=50, num_attackers=55, terrain_parameters=[0.15, 0.2], REWARD_DEBUG=True, REWARD_DEBUG_Value=3),
}


def bytes_per_env_step(G, A):
    """obs G*768 + state r/w 2*N*60 + actions G*4 + attacker force A*16 + reward G*4 +
    done/win 2 + alive N."""
    N = G + A
    return G * 768 + 2 * N * 60 + G * 4 + A * 16 + G * 4 + 2 + N
# SURVEY.md section 8(d), gather kernel (k_observe): obs G*768 + state read N*57
GATHER_BYTES_PER_ENV = 50 * 768 + 100 * 57
FALLBACK_HBM_GBS = 6650.0


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per k_step launch from the committed ncu --set full capture."""
    p = os.path.join(REPO, 'profiles', 'latest_ncu.json')
    try:
        return json.load(open(p))
    except Exception:
        return {}


def dram_frac(ncu, kernel, kernel_ms, peak, E, scenario):
    """DRAM-counter roofline fraction: dram__bytes (read + write) per launch, captured by ncu over a RANGE of
    back-to-back launches (tools/dram_range.py -> profiles/latest_ncu.json, so write-backs deferred past a kernel's end
    are charged), divided by the live CUDA-event duration and the measured peak.  None without a matching capture."""
    b = ncu.get(kernel + '_dram_bytes_range_per_launch')
    if not b or E != ncu.get('envs_per_gpu') or scenario != '50v50':
        return None
    return b / (kernel_ms * 1e-3) / 1e9 / peak


def measured_peak():
    p = os.path.join(REPO, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        try:
            return float(json.load(open(p))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
        except Exception:
            pass
    return FALLBACK_HBM_GBS, 'fallback (B200_PROFILING.md 6.65 TB/s)'


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.01):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.sm_max = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: 'hw_slowdown',
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: 'hw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: 'sw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonSwPowerCap: 'sw_power_cap',
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: 'hw_power_brake'}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {'sm_mhz': med, 'sm_max_mhz': self.sm_max, 'reasons': sorted(self.reasons), 'samples': len(self.samples)}


def physical_gpu_index(local_index):
    vis = os.environ.get('CUDA_VISIBLE_DEVICES')
    if vis:
        try:
            return int(vis.split(',')[local_index])
        except Exception:
            return local_index
    return local_index


# ------------------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_port_run(n_env, steps, warmup, threads, seed=SEED, full_planner=True):
    """Times the CPU restatement of the reference path (oracle/dca_oracle.c + dca_planner.c, OpenMP over envs) on
    `threads` host threads: scripted attacker planner + step + obs + autoreset, same workload (what one
    SuperpoolEnv.step does in the reference).  Returns (agent_steps_per_s, seconds)."""
    sys.path.insert(0, os.path.join(REPO, 'oracle'))
    from dca_oracle import OraclePlanner, OracleVecEnv
    env = OracleVecEnv(WORKLOAD, n_env, seed=seed, nthreads=threads)
    env.reset()
    planner = OraclePlanner(env)
    plan = planner.plan if full_planner else env.plan_centroid
    rng = np.random.RandomState(1)
    pool = [rng.randint(0, 7, size=(n_env, env.G)).astype(np.float64) for _ in range(8)]
    for i in range(warmup):
        env.step(pool[i % 8], plan())
    t0 = time.perf_counter()
    for i in range(steps):
        env.step(pool[i % 8], plan())
    dt = time.perf_counter() - t0
    return n_env * env.N * steps / dt, dt


def reference_pool_run(scenario, steps, warmup, procs=None, fold=2, timeout=600):
    """The UNMODIFIED reference path (SmartPool + SuperpoolEnv + EnvAutoReset + MISSION/dca, staged into oracle/_ref by
    __graft_entry__.build()) timed in its own process by oracle/ref_pool_bench.py.  None if it is not staged or fails."""
    import subprocess
    sys.path.insert(0, os.path.join(REPO, 'oracle'))
    import stage_ref
    if not stage_ref.staged():
        return None
    procs = procs or (os.cpu_count() or 1)
    cmd = [sys.executable, os.path.join(REPO, 'oracle', 'ref_pool_bench.py'), '--scenario', scenario, '--procs', str(procs),
           '--fold', str(fold), '--steps', str(steps), '--warmup', str(warmup), '--seed', str(SEED)]
    try:
        out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=timeout, text=True).stdout
    except Exception as e:  # pragma: no cover
        sys.stderr.write('reference pool run failed: %r\n' % (e,))
        return None
    for ln in out.splitlines():
        if ln.startswith('REF_POOL_JSON '):
            return json.loads(ln[len('REF_POOL_JSON '):])
    sys.stderr.write('reference pool run printed no result:\n' + out[-2000:] + '\n')
    return None


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores -- SmartPool with
    one worker process per host core, fold 2 (E_cpu = 2 x cores), SuperpoolEnv.step with random guard
    actions; the scripted attacker planner runs inside World.step as it always does in the reference.  Each step is a
    bounded sample (2 x cores envs instead of 4096 per GPU: env-steps/s of the process pool does not depend on how many
    envs queue behind each worker).  If oracle/_ref is not staged the arm falls back to the C/OpenMP port of the same
    work and says so (kind: "port")."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    steps = max(1, min(args.steps, 200))
    warm = max(3, min(args.warmup, 20))
    ref = reference_pool_run(args.scenario, steps, warm, procs=cores, fold=2)
    # the C/OpenMP port of the same work on the same cores: the port / reference calibration on ONE machine
    port_v, port_dt = cpu_port_run(1024, min(steps, 60), 5, cores)
    port_v0, _ = cpu_port_run(1024, min(steps, 60), 5, cores, full_planner=False)
    if ref is not None:
        value, ms, kind = ref['agent_steps_per_sec'], ref['ms_per_step'], 'reference'
        sample = ('%d envs (SmartPool: %d worker processes x fold %d) x %d steps of the %s workload through '
                  'SuperpoolEnv.step, unmodified reference staged in oracle/_ref (%.1f s)'
                  % (ref['envs'], ref['procs'], ref['fold'], steps, args.scenario, ref['seconds']))
        n_env = ref['envs']
    else:
        value, ms, kind = port_v, port_dt / min(steps, 60) * 1e3, 'port'
        sample = '1024 envs x %d steps incl. the scripted attacker planner, C port, OpenMP over envs' % min(steps, 60)
        n_env = 1024
    line = {
        'impl': 'reference', 'metric': 'agent_steps_per_sec', 'value': value, 'unit': 'agent-steps/s',
        'n_gpus': args.gpus, 'steps': steps, 'warmup': warm, 'ms_per_step': ms,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': 'DCA %s scripted-attacker planner + step + obs + autoreset, injected random guard actions'
                               % args.scenario,
                   'envs_per_step_sample': n_env, 'scenario': WORKLOAD},
        'cpu_baseline': {'value': value, 'unit': 'agent-steps/s', 'cores': cores, 'kind': kind, 'sample': sample,
                         'port_value': port_v, 'port_value_without_planner': port_v0,
                         'port_over_reference': (port_v / value) if kind == 'reference' else None,
                         'note': 'port_* = oracle/dca_oracle.c + dca_planner.c (OpenMP, same cores, same workload); the '
                                 'reference cannot run without its planner (core.py:189-196), so value_without_planner '
                                 'exists for the port only'},
        'value_without_planner': port_v0 if kind == 'port' else None,
        'e2e': {'value': value, 'unit': 'agent-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ GPU arm
def policy_rollout_run(E, dev, rank, world, steps=60, warm=10):
    """BASELINE.json configs[3] on the device end to end: RESULT/50RL-55opp -- 50 guards
    driven by the shipped conc_4hist policy (sampled actions) against 55 scripted attackers: k_histpool -> DynamicNorm
    kernels + torch Net -> k_plan + k_step -> rollout buffer, then k_gae + advantage normalisation over the buffer.
    Nothing crosses PCIe inside the loop.  The checkpoint comes from tests/golden/policy_net_4hist.npz (the state dict
    of RESULT/50RL-55opp/model.pt as arrays, written by oracle/make_golden_policy.py)."""
    import torch
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.policy import ConcNet4Hist, PolicyRollout
    from hmp2g_b200.vec_env import DcaVecEnv
    path = os.path.join(REPO, 'tests', 'golden', 'policy_net_4hist.npz')
    if not os.path.exists(path):
        return None
    z = np.load(path)
    net = ConcNet4Hist(rawob_dim=16, n_action=7, n_focus_on=3)
    net.load_state_dict({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith('sd/')})
    net = net.to(dev).train()
    wl = dict(WORKLOAD, **SCENARIOS['50v55'])
    scfg = type('BenchScenarioConfig_policy', (ScenarioConfig,), wl)
    env = DcaVecEnv(E, scfg=scfg, device=dev, seed=SEED, env_gid0=rank * E, empty_rows_nan=True)
    roll = PolicyRollout(env, net, gamma=0.99, tau=0.95, gamma_fwd=0.95, seed=SEED + rank)
    stream = torch.cuda.current_stream(dev)
    for _ in range(warm):
        roll.step()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0.record(stream)
    for _ in range(steps):
        roll.step()
    t1.record(stream)
    torch.cuda.synchronize()
    ms_eager = t0.elapsed_time(t1) / steps
    replay = roll.capture(warm=2)   # the same step as one CUDA graph launch
    for _ in range(3):
        replay()
    torch.cuda.synchronize()
    t0.record(stream)
    for _ in range(steps):
        replay()
    t1.record(stream)
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / steps
    # the policy forward alone (history roll + normalisation + Net.act) on the last observation
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(stream)
    for _ in range(5):
        feed = roll.hist.roll(roll.obs, suffered_reset=roll.suffered)
        net.act(feed, generator=roll.gen)
    p1.record(stream)
    torch.cuda.synchronize()
    ms_policy = p0.elapsed_time(p1) / 5
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(stream)
    ret, threat, adv, valid = roll.buf.finalize(None)   # every env's (truncated) trajectory: GAE + normalised advantage
    g1.record(stream)
    torch.cuda.synchronize()
    out = {'workload': 'RESULT/50RL-55opp: DCA 50 guards (conc_4hist policy, shipped model.pt, sampled actions) vs 55 '
                       'scripted attackers (k_plan), %d envs/GPU, device-resident loop' % E,
           'value': world * E * env.N / (ms * 1e-3), 'unit': 'agent-steps/s', 'ms_per_step': ms,
           'ms_per_step_eager': ms_eager, 'launch': 'one CUDA graph per env step (PolicyRollout.capture)',
           'policy_forward_ms_eager': ms_policy, 'steps': steps,
           'gae_and_advantage_ms': g0.elapsed_time(g1), 'rollout_buffer_TEG': list(roll.buf.reward.shape),
           'episodes_finished': int(roll.episodes.item()), 'wins': int(roll.wins.item()),
           'policy': 'ConcNet4Hist (hmp2g_b200/policy.py): encoder + normalisation (k_policy_encode), attention / top-k '
                     '(k_conc_attend), every dense layer behind them (dca_policy_linear, 3xTF32 on the tensor cores), '
                     'moments, history roll, planner, step, GAE: this library; torch: two 48 x 48 weight folds, the '
                     'Categorical draw and the rollout-buffer bookkeeping',
           'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    del roll, env, net
    torch.cuda.empty_cache()
    return out


def quick_config_run(name, E, dev, rank, steps=360, warm=40):
    """BASELINE.json configs[2] / [3] (100v100, 50v55) at the same env count: step+obs+autoreset with the centroid
    stand-in (value) and k_step alone per launch over two episodes (kernel_ms, roofline fraction), plus k_plan."""
    import ctypes as C
    import torch
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.vec_env import DcaVecEnv
    wl = dict(WORKLOAD, **SCENARIOS[name])
    scfg = type('BenchScenarioConfig_' + name, (ScenarioConfig,), wl)
    env = DcaVecEnv(E, scfg=scfg, device=dev, seed=SEED, env_gid0=rank * E)
    G, A, N = env.G, env.A, env.N
    lib, h = env.lib, env.handle
    stream = torch.cuda.current_stream(dev)
    sptr = C.c_void_p(stream.cuda_stream)
    force = torch.zeros((E, A, 2), dtype=torch.float64, device=dev)
    fptr = C.c_void_p(force.data_ptr())
    pool = [torch.randint(0, 7, (E, G), device=dev, dtype=torch.int32) for _ in range(8)]
    aptrs = [C.c_void_p(a.data_ptr()) for a in pool]
    env.reset()
    lib.dca_set_option(h, 4, 1)  # value: launches pipelined as in the headline run; k_step_ms below: stand-alone
    for i in range(warm):
        lib.dca_plan_centroid(h, fptr, sptr)
        lib.dca_step(h, aptrs[i % 8], fptr, None, None, 1, sptr)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0.record(stream)
    for i in range(steps):
        lib.dca_plan_centroid(h, fptr, sptr)
        lib.dca_step(h, aptrs[i % 8], fptr, None, None, 1, sptr)
    t1.record(stream)
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / steps
    # the same as two env groups on their own streams (DcaEnvGroups), as in the headline run
    from hmp2g_b200.vec_env import DcaEnvGroups
    grp = DcaEnvGroups(E, groups=2, scfg=scfg, device=dev, seed=SEED, env_gid0=rank * E)
    grp.reset()
    for i in range(warm):
        grp.step(pool[i % 8], planner='centroid', sync=False)
    torch.cuda.synchronize()
    t0.record(stream)
    grp.fork()
    for i in range(steps):
        grp.step(pool[i % 8], planner='centroid', sync=False)
    grp.join()
    t1.record(stream)
    torch.cuda.synchronize()
    ms_groups = t0.elapsed_time(t1) / steps
    grp.close()
    del grp
    es = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ee = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    lib.dca_set_option(h, 2, 0)
    lib.dca_set_option(h, 4, 0)
    for i in range(steps):
        lib.dca_plan_centroid(h, fptr, sptr)
        es[i].record(stream)
        lib.dca_step(h, aptrs[i % 8], fptr, None, None, 1, sptr)
        ee[i].record(stream)
    torch.cuda.synchronize()
    k_ms = sum(es[i].elapsed_time(ee[i]) for i in range(steps)) / steps
    ps = [torch.cuda.Event(enable_timing=True) for _ in range(180)]
    pe = [torch.cuda.Event(enable_timing=True) for _ in range(180)]
    for i in range(180):
        ps[i].record(stream)
        lib.dca_plan(h, fptr, None, None, sptr)
        pe[i].record(stream)
        lib.dca_step(h, aptrs[i % 8], fptr, None, None, 1, sptr)
    torch.cuda.synchronize()
    plan_ms = sum(ps[i].elapsed_time(pe[i]) for i in range(180)) / 180
    peak, _ = measured_peak()
    algo = bytes_per_env_step(G, A) * E
    out = {'workload': 'DCA %dv%d step+obs+autoreset, %d envs, centroid stand-in' % (G, A, E),
           'value': E * N / (min(ms, ms_groups) * 1e-3), 'unit': 'agent-steps/s', 'ms_per_step': min(ms, ms_groups),
           'ms_per_step_one_group_pipelined': ms, 'ms_per_step_two_groups': ms_groups,
           'k_step_ms': k_ms, 'k_plan_ms': plan_ms,
           'algorithmic_bytes_per_launch': algo, 'achieved_gbs': algo / (k_ms * 1e-3) / 1e9,
           'frac': algo / (k_ms * 1e-3) / 1e9 / peak, 'steps': steps, 'threads_per_cta': 128 if N <= 128 else 256}
    env.close()
    del env, force, pool
    torch.cuda.empty_cache()
    return out


def collective_timings(E, G, dev, world):
    """the three collectives of a sharded rollout, on NCCL: (i) flat fp32 gradient all-reduce of the conc
    policy (63 370 parameters), (ii) advantage moments inside normalized_advantage over a [180,E,G] buffer, (iii) rollout
    statistics.  CUDA events, max over ranks, microseconds per call."""
    import torch
    import torch.distributed as dist
    from hmp2g_b200.gae import allreduce_moments, normalized_advantage
    from hmp2g_b200.sharding import RolloutStats
    stream = torch.cuda.current_stream(dev)

    def timed(fn, n):
        for _ in range(3):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        dist.barrier()
        a.record(stream)
        for _ in range(n):
            fn()
        b.record(stream)
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / n * 1e3], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    grad = torch.randn(63370, device=dev)
    T = 180 if E * G * 180 * 12 < 4e9 else 45
    ret = torch.randn((T, E, G), device=dev)
    val = torch.randn((T, E, G), device=dev)
    stats3 = torch.ones(3, dtype=torch.float64, device=dev)
    rs = RolloutStats(dev)
    out = {'world': world, 'backend': dist.get_backend(),
           'grad_allreduce_us': timed(lambda: dist.all_reduce(grad), 50), 'grad_params': 63370,
           'moments_allreduce_us': timed(lambda: allreduce_moments(stats3), 50),
           'normalized_advantage_us': timed(lambda: normalized_advantage(ret, val), 10), 'advantage_shape': [T, E, G],
           'rollout_stats_reduce_us': timed(lambda: rs.reduce(), 20)}
    del ret, val
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.sharding import init_distributed
    from hmp2g_b200.vec_env import DcaSuperpoolEnv, DcaVecEnv

    assert torch.cuda.is_available(), 'bench.py needs a GPU: the product path has no CPU fallback'
    rank, world, local_rank = init_distributed('nccl')
    assert world == args.gpus, 'launch with torchrun --nproc-per-node %d (WORLD_SIZE=%d)' % (args.gpus, world)
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    E, K, W = args.envs_per_gpu, args.steps, args.warmup
    scfg = type('BenchScenarioConfig', (ScenarioConfig,), dict(WORKLOAD))
    env = DcaVecEnv(E, scfg=scfg, device=dev, seed=SEED, env_gid0=rank * E)
    G, A, N = env.G, env.A, env.N
    gen = torch.Generator(device=dev)
    gen.manual_seed(SEED + rank)
    n_pool = 16  # pre-generated guard action batches resident in HBM (RandomController, ALGORITHM/random/foundation.py:28)
    act_pool = [torch.randint(0, 7, (E, G), generator=gen, device=dev, dtype=torch.int32) for _ in range(n_pool)]
    force = torch.zeros((E, A, 2), dtype=torch.float64, device=dev)
    lib, h = env.lib, env.handle
    import ctypes as C
    stream = torch.cuda.current_stream(dev)
    sptr = C.c_void_p(stream.cuda_stream)
    fptr = C.c_void_p(force.data_ptr())
    aptrs = [C.c_void_p(a.data_ptr()) for a in act_pool]

    if args.planner == 'scripted':  # the reference's formation planner on the GPU (k_plan) instead of the stand-in
        def plan_call(h_, f_, s_):
            return lib.dca_plan(h_, f_, None, None, s_)
    else:
        plan_call = lib.dca_plan_centroid

    def one_step(i):
        rc = plan_call(h, fptr, sptr)
        rc |= lib.dca_step(h, aptrs[i % n_pool], fptr, None, None, 1, sptr)
        return rc

    overlap = 0 if args.no_overlap else 1
    piped = 0 if (args.no_pipeline or args.no_overlap) else 1
    lib.dca_set_option(h, 2, overlap)
    lib.dca_set_option(h, 4, piped)  # DCA_OPT_PIPELINE_STEPS: plan / step / plan / step hand envs over one by one
    env.reset()
    # Pre-roll (set-up, untimed): env e is advanced by a random number of steps in [0, MaxEpisodeStep) -- it sits out
    # (ENV-PAUSE mask) the first MaxEpisodeStep - offset[e] of MaxEpisodeStep steps -- so the timed steps sample the
    # stationary mix of episode phases of a long-running job instead of steps W..W+K of 4096 synchronised first
    # episodes (a driver run of 20 steps would otherwise sit entirely in the most expensive phase, where the teams
    # meet).  --no-preroll times the synchronised start.
    preroll = 0 if args.no_preroll else int(WORKLOAD.get('MaxEpisodeStep', 180))
    if preroll:
        offs = torch.randint(0, preroll, (E,), generator=gen, device=dev)
        for s in range(preroll):
            skip = (offs < (preroll - s)).to(torch.uint8)
            plan_call(h, fptr, sptr)
            assert lib.dca_step(h, aptrs[s % n_pool], fptr, C.c_void_p(skip.data_ptr()), None, 1, sptr) == 0
        torch.cuda.synchronize()
    sampler = ClockSampler(physical_gpu_index(local_rank), period=0.002) if rank == 0 else None
    # ---- the headline region with --env-groups n > 1: the E envs as n independent groups (DcaEnvGroups: E / n envs
    # each, contiguous global env ids, own streams; same results as one group, tests/test_gpu_parity.py) -- one group's
    # CTAs fill the slots the other group's launch boundary leaves empty.  Per-env pipelining is off in the groups.
    # (auto: two groups from one full wave of CTAs per group upwards -- profiles/r2_cn_E_sweep_groups.txt)
    n_groups = args.env_groups if args.env_groups > 0 else (2 if args.planner == 'centroid' and E >= 4096 else 1)
    grouped = None
    if n_groups > 1:
        from hmp2g_b200.vec_env import DcaEnvGroups
        grp = DcaEnvGroups(E, groups=n_groups, scfg=scfg, device=dev, seed=SEED, env_gid0=rank * E,
                           overlap_planner=bool(overlap))
        grp.reset()
        Eg = grp.Eg
        g_s = [C.c_void_p(s.cuda_stream) for s in grp.streams]
        g_f = [C.c_void_p(ge._force.data_ptr()) for ge in grp.envs]
        g_a = [[C.c_void_p(a[k * Eg:(k + 1) * Eg].data_ptr()) for a in act_pool] for k in range(n_groups)]

        def group_step(i, skips=None):
            for k, ge in enumerate(grp.envs):
                plan_call(ge.handle, g_f[k], g_s[k])
                lib.dca_step(ge.handle, g_a[k][i % n_pool], g_f[k],
                             None if skips is None else C.c_void_p(skips.data_ptr() + k * Eg), None, 1, g_s[k])
        if preroll:
            all_skips = (offs[None, :] < (preroll - torch.arange(preroll, device=dev))[:, None]).to(torch.uint8)
            torch.cuda.synchronize()
            for s in range(preroll):
                group_step(s, all_skips[s])
            torch.cuda.synchronize()
            del all_skips
        for i in range(W):
            group_step(i)
        gt0, gt1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if sampler:
            sampler.start()
        gl0 = grp.kernel_launches()
        gt0.record(stream)
        for s_ in grp.streams:
            s_.wait_event(gt0)
        for i in range(K):
            group_step(W + i)
        grp.join()
        gt1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        tg = torch.tensor([gt0.elapsed_time(gt1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tg, op=dist.ReduceOp.MAX)
        grouped = {'ms_total': float(tg.item()), 'gpu_launches': grp.kernel_launches() - gl0,
                   'alive_frac': float(grp.alive.float().mean().item()), 'done_frac': float(grp.done.float().mean().item())}
        grp.close()
        del grp
        torch.cuda.empty_cache()
    launches0 = env.kernel_launches()
    for i in range(W):
        assert one_step(i) == 0
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if sampler and grouped is None:
        sampler.start()
    launches_before = env.kernel_launches()
    # timed region: K steps back to back, nothing else on the stream (the step kernel is a programmatic dependent of
    # the planner stand-in, DCA_OPT_OVERLAP_PLANNER: an event between the two launches would serialise them)
    t_begin.record(stream)
    for i in range(K):
        plan_call(h, fptr, sptr)
        lib.dca_step(h, aptrs[(W + i) % n_pool], fptr, None, None, 1, sptr)
    t_end.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    gpu_launches = env.kernel_launches() - launches_before
    ms_total = t_begin.elapsed_time(t_end)
    # k_step's own launch duration (the roofline figure): the next Kk steps of the same episodes, the kernel launched
    # stand-alone (no overlap with the planner stand-in) between two events each
    Kk = min(max(K, 360), 720)  # at least two whole episodes, so a short driver run also reports the episode average
    Kfirst = min(K, Kk)
    ev_s = [torch.cuda.Event(enable_timing=True) for _ in range(Kk)]
    ev_e = [torch.cuda.Event(enable_timing=True) for _ in range(Kk)]
    lib.dca_set_option(h, 2, 0)
    lib.dca_set_option(h, 4, 0)
    for i in range(Kk):
        plan_call(h, fptr, sptr)
        ev_s[i].record(stream)
        lib.dca_step(h, aptrs[(W + K + i) % n_pool], fptr, None, None, 1, sptr)
        ev_e[i].record(stream)
    torch.cuda.synchronize()
    lib.dca_set_option(h, 2, overlap)
    lib.dca_set_option(h, 4, piped)
    per_launch = [ev_s[i].elapsed_time(ev_e[i]) for i in range(Kk)]
    ms_step_kernel = sum(per_launch[:Kfirst]) / Kfirst  # the K steps that follow the timed region
    ms_step_kernel_episode = sum(per_launch) / Kk       # >= two whole episodes
    # ---- like-for-like with the reference arm: the scripted formation planner (k_plan, what World.step runs in the
    # reference, core.py:189-196) + step, over whole episodes, the step launched under its planner as in the timed
    # region.  Its own env instance, driven by the scripted planner from the reset on (states produced by the centroid
    # stand-in -- attackers collapsed onto one point -- are not states this planner ever sees), same pre-roll.
    env2 = DcaVecEnv(E, scfg=scfg, device=dev, seed=SEED + 1, env_gid0=rank * E)
    h2 = env2.handle
    lib.dca_set_option(h2, 2, overlap)
    lib.dca_set_option(h2, 4, piped)
    env2.reset()
    if preroll:
        for s in range(preroll):
            skip = (offs < (preroll - s)).to(torch.uint8)
            lib.dca_plan(h2, fptr, None, None, sptr)
            lib.dca_step(h2, aptrs[s % n_pool], fptr, C.c_void_p(skip.data_ptr()), None, 1, sptr)
    Ks = 360
    for i in range(20):
        lib.dca_plan(h2, fptr, None, None, sptr)
        lib.dca_step(h2, aptrs[i % n_pool], fptr, None, None, 1, sptr)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    s0.record(stream)
    for i in range(Ks):
        lib.dca_plan(h2, fptr, None, None, sptr)
        lib.dca_step(h2, aptrs[(20 + i) % n_pool], fptr, None, None, 1, sptr)
    s1.record(stream)
    torch.cuda.synchronize()
    ms_scripted = s0.elapsed_time(s1) / Ks
    pe_s = [torch.cuda.Event(enable_timing=True) for _ in range(180)]
    pe_e = [torch.cuda.Event(enable_timing=True) for _ in range(180)]
    lib.dca_set_option(h2, 2, 0)
    lib.dca_set_option(h2, 4, 0)
    for i in range(180):  # k_plan alone, on the states of one more episode
        pe_s[i].record(stream)
        lib.dca_plan(h2, fptr, None, None, sptr)
        pe_e[i].record(stream)
        lib.dca_step(h2, aptrs[i % n_pool], fptr, None, None, 1, sptr)
    torch.cuda.synchronize()
    ms_plan_episode = sum(pe_s[i].elapsed_time(pe_e[i]) for i in range(180)) / 180
    del env2
    ts_ = torch.tensor([ms_scripted], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ts_, op=dist.ReduceOp.MAX)
    scripted = {'value': world * E * N / (float(ts_.item()) * 1e-3), 'unit': 'agent-steps/s',
                'ms_per_step': float(ts_.item()), 'k_plan_ms': ms_plan_episode, 'steps': Ks,
                'what': 'dca_plan (scripted formation planner, cheat_script_ai.py:28-194) + dca_step per step over two '
                        'episodes: the work of one World.step + observation in the reference (--impl reference times '
                        'exactly this on the host)'}
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * E * N * K / (ms_total * 1e-3)
    alive_frac = float(env.alive.float().mean().item())
    done_frac = float(env.done.float().mean().item())
    one_group = {'value': value, 'unit': 'agent-steps/s', 'ms_per_step': ms_total / K, 'gpu_launches': gpu_launches,
                 'pipeline_steps': bool(piped),
                 'what': 'the same K steps with all %d envs in one DcaVecEnv on one stream (launch t+1 handed its envs '
                         'one by one, DCA_OPT_PIPELINE_STEPS)' % E}
    if grouped is not None:  # the headline is the grouped region
        ms_total, gpu_launches = grouped['ms_total'], grouped['gpu_launches']
        value = world * E * N * K / (ms_total * 1e-3)
        alive_frac, done_frac = grouped['alive_frac'], grouped['done_frac']

    # ---- the gather kernel alone (k_observe: kNN + obs gather on the resident state), north_star's ">= 60 % of HBM
    # roofline on the gather kernel" line
    Kg = 200
    for _ in range(5):
        lib.dca_observe(h, None, sptr)
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(stream)
    for _ in range(Kg):
        lib.dca_observe(h, None, sptr)
    g1.record(stream)
    torch.cuda.synchronize()
    ms_gather = g0.elapsed_time(g1) / Kg
    # ... and as two env groups on their own streams (the headline's arrangement: one group's launch boundary is covered
    # by the other group's CTAs), on the observation state of a short rollout of their own
    from hmp2g_b200.vec_env import DcaEnvGroups
    ggrp = DcaEnvGroups(E, groups=2, scfg=scfg, device=dev, seed=SEED, env_gid0=rank * E)
    ggrp.reset()
    for i in range(60):
        ggrp.step(act_pool[i % n_pool], planner='centroid', sync=False)
    for _ in range(5):
        ggrp.observe(sync=False)
    torch.cuda.synchronize()
    g0.record(stream)
    ggrp.fork()
    for _ in range(Kg):
        ggrp.observe(sync=False)
    ggrp.join()
    g1.record(stream)
    torch.cuda.synchronize()
    ms_gather_groups = g0.elapsed_time(g1) / Kg
    ggrp.close()
    del ggrp
    # ---- GAE / return kernel over a rollout buffer [T,E,G] (21 B per (t, env, guard))
    from hmp2g_b200.gae import gae_return
    T = 180 if E * G * 180 * 21 < 8e9 else 45
    r_buf = torch.rand((T, E, G), device=dev)
    v_buf = torch.rand((T, E, G), device=dev)
    d_buf = (torch.rand((T, E, G), device=dev) < 0.2).to(torch.uint8)
    for _ in range(2):
        gae_return(r_buf, v_buf, d_buf)
    g0.record(stream)
    for _ in range(5):
        gae_return(r_buf, v_buf, d_buf)
    g1.record(stream)
    torch.cuda.synchronize()
    ms_gae = g0.elapsed_time(g1) / 5
    gae_bytes = T * E * G * (4 + 4 + 1 + 4 + 4)  # reward, value, dead in; return, threat out
    del r_buf, v_buf, d_buf
    # ---- conc_4hist history pool (k_histpool) on the env's observations + runner bookkeeping (k_runner_*)
    from hmp2g_b200.rollout import HistoryPool
    from hmp2g_b200.runner import DeviceRunner
    hp = HistoryPool(E, G, device=dev)
    run = DeviceRunner(E, G, align_episode=False, report_interval=64, device=dev)
    for _ in range(5):
        hp.roll(env.obs)
        run.update(env.done, env.win, env.reward)
    g0.record(stream)
    for _ in range(50):
        hp.roll(env.obs)
    g1.record(stream)
    torch.cuda.synchronize()
    ms_hist = g0.elapsed_time(g1) / 50
    g0.record(stream)
    for _ in range(50):
        run.update(env.done, env.win, env.reward)
    g1.record(stream)
    torch.cuda.synchronize()
    ms_runner = g0.elapsed_time(g1) / 50
    hist_bytes = E * G * 12 * 16 * 4 * 5  # obs + history in, obs_feed (2x) + history out
    del hp, run
    # ---- dca_multiteam (k_mt_step): 2 teams x 25, every agent observed and acting, same env count
    from hmp2g_b200.multiteam import DcaMultiTeamVecEnv, MtScenarioConfig
    mt_cfg = type('S', (MtScenarioConfig,), dict(N_AGENT_EACH_TEAM=[25, 25], random_jam_prob=0.05, introduce_terrain=True,
                                                 terrain_parameters=[0.05, 0.2]))
    mt = DcaMultiTeamVecEnv(E, scfg=mt_cfg, device=dev, seed=SEED, env_gid0=rank * E)
    mt.reset()
    mt_pool = [torch.randint(0, 7, (E, mt.N), device=dev, dtype=torch.int32) for _ in range(8)]
    for i in range(20):
        mt.step(mt_pool[i % 8])
    Km = 300
    g0.record(stream)
    for i in range(Km):
        mt.step(mt_pool[i % 8])
    g1.record(stream)
    torch.cuda.synchronize()
    ms_mt = g0.elapsed_time(g1) / Km
    # obs N*816 out; state (pos, vel, atk_rad, terrain f64 + 5 flag bytes) read + written; actions in; reward + alive out
    mt_bytes = E * mt.N * (12 * 17 * 4 + 2 * (6 * 8 + 5) + 4 + 4 + 1)
    mt_alive = float(mt.alive.float().mean().item())
    mt_N = mt.N
    del mt
    # the same envs as two groups of E / 2 on their own streams (contiguous global env ids: the same episodes)
    mt_streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    mt_g, mt_gp = [], []
    for k in range(2):
        with torch.cuda.stream(mt_streams[k]):
            m_ = DcaMultiTeamVecEnv(E // 2, scfg=mt_cfg, device=dev, seed=SEED, env_gid0=rank * E + k * (E // 2))
            m_.reset()
            mt_g.append(m_)
            mt_gp.append([p[k * (E // 2):(k + 1) * (E // 2)].contiguous() for p in mt_pool])
    torch.cuda.synchronize()

    def mt_group_step(i):
        for k in range(2):
            with torch.cuda.stream(mt_streams[k]):
                mt_g[k].step(mt_gp[k][i % 8])
    for i in range(20):
        mt_group_step(i)
    torch.cuda.synchronize()
    g0.record(stream)
    for s_ in mt_streams:
        s_.wait_event(g0)
    for i in range(Km):
        mt_group_step(i)
    for s_ in mt_streams:
        ev_ = torch.cuda.Event()
        ev_.record(s_)
        stream.wait_event(ev_)
    g1.record(stream)
    torch.cuda.synchronize()
    ms_mt_groups = g0.elapsed_time(g1) / Km
    del mt_g, mt_gp, mt_pool
    # ---- BASELINE.json configs[2] / [3] at the same env count (extra report lines; not the headline)
    other = {}
    if args.scenario == '50v50' and not args.no_other_configs:
        for name in ('100v100', '50v55'):
            other[name] = quick_config_run(name, E, dev, rank)
    coll = collective_timings(E, G, dev, world) if world > 1 else None
    policy_line = None
    if args.scenario == '50v50' and not args.no_other_configs:
        policy_line = policy_rollout_run(E, dev, rank, world)
    # ---- k_step from a synchronised reset (what a 20-step run measured before the pre-roll existed): steps 5..25 of
    # 4096 first episodes in lock-step, the most expensive phase (the teams meet)
    env.reset()
    lib.dca_set_option(h, 2, 0)
    lib.dca_set_option(h, 4, 0)
    for i in range(5):
        plan_call(h, fptr, sptr)
        lib.dca_step(h, aptrs[i % n_pool], fptr, None, None, 1, sptr)
    ss = [torch.cuda.Event(enable_timing=True) for _ in range(20)]
    se = [torch.cuda.Event(enable_timing=True) for _ in range(20)]
    for i in range(20):
        plan_call(h, fptr, sptr)
        ss[i].record(stream)
        lib.dca_step(h, aptrs[(5 + i) % n_pool], fptr, None, None, 1, sptr)
        se[i].record(stream)
    torch.cuda.synchronize()
    ms_step_sync_start = sum(ss[i].elapsed_time(se[i]) for i in range(20)) / 20

    # ---- e2e: the reference-facing call (SuperpoolEnv.step surface) with HOST buffers, copies inside the timed region
    del env
    torch.cuda.empty_cache()
    from hmp2g_b200.config import GlobalConfig
    GlobalConfig.seed = SEED
    host = DcaSuperpoolEnv(n_env=E, scfg=scfg, device=dev, env_gid0=rank * E)
    host.reset()
    rng = np.random.RandomState(rank)
    host_pool = [rng.randint(0, 7, size=(E, G)).astype(np.float64) for _ in range(4)]
    Ke = max(3, min(K, args.e2e_steps))
    for i in range(3):
        host.step(host_pool[i % 4])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(Ke):
        host.step(host_pool[i % 4])
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * E * N * Ke / float(te.item())
    h2d, d2h = host.h2d_bytes, host.d2h_bytes
    # ---- the scripted attacker planner kernel (k_plan) on the state its own rollout produced; part of e2e, not of
    # configs[1] ("injected actions")
    pf = C.c_void_p(host.env._force.data_ptr())
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(stream)
    for _ in range(20):
        host.env.lib.dca_plan(host.env.handle, pf, None, None, sptr)
    g1.record(stream)
    torch.cuda.synchronize()
    ms_plan = g0.elapsed_time(g1) / 20

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peak()
    algo_bytes = bytes_per_env_step(G, A) * E
    gather_bytes = (G * 768 + N * 57) * E
    ncu = ncu_traffic()
    traffic = args.traffic if args.traffic is not None else (
        ncu.get('k_step_dram_bytes_per_launch') if (E == ncu.get('envs_per_gpu') and args.scenario == '50v50') else None)
    achieved_alone = algo_bytes / (ms_step_kernel * 1e-3) / 1e9
    # roofline.achieved: the algorithmic bytes of the k_step launches of the timed region / the region's time (CUDA events
    # on the launch stream).  Launches overlap there (env groups / per-env hand-over), so this is the only launch
    # duration the region defines; the stand-alone duration of one launch over all E envs is reported next to it.
    achieved = algo_bytes / (ms_total / K * 1e-3) / 1e9
    line = {
        'metric': 'agent_steps_per_sec', 'value': value, 'unit': 'agent-steps/s', 'n_gpus': world, 'steps': K,
        'warmup': W, 'ms_per_step': ms_total / K, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': 'DCA %dv%d step+obs+autoreset, %d envs/GPU, injected random guard actions, '
                               '%s (BASELINE.json configs[1]%s)'
                               % (G, A, E, 'centroid planner stand-in' if args.planner == 'centroid' else
                                  'scripted formation planner k_plan in the timed region',
                                  '' if args.scenario == '50v50' else ' at another team size'),
                   'envs_per_gpu': E, 'entities_per_env': N, 'scenario': WORKLOAD, 'overlap_planner': bool(overlap),
                   'env_groups': n_groups,
                   'env_groups_note': ('the %d envs step as %d independent groups of %d (DcaEnvGroups: own streams, '
                                       'contiguous global env ids, results identical to one group)' % (E, n_groups, E // n_groups))
                   if grouped is not None else 'one group',
                   'pipeline_steps': bool(piped) if grouped is None else False,
                   'pre_roll': ('each env advanced by a random number of steps in [0, %d) before warm-up (untimed set-up): '
                                'stationary mix of episode phases' % preroll) if preroll else 'none: synchronised reset',
                   'l2': 'no explicit flush: per-step working set %.0f MB (obs %.0f MB written + state %.0f MB r/w) '
                         '> 126 MB L2' % ((E * (G * 768 + 2 * 5760)) / 1e6, E * G * 768 / 1e6, 2 * E * 5760 / 1e6),
                   'policy_agent_steps_per_sec': value * G / N, 'planner_kernel_ms': ms_plan,
                   'alive_fraction_at_end': alive_frac, 'done_fraction_last_step': done_frac},
        'roofline': {'bound': 'hbm', 'kernel': 'k_step', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                     'frac': achieved / peak, 'traffic': traffic, 'peak_source': peak_src,
                     'algorithmic_bytes_per_launch': algo_bytes,
                     'achieved_source': 'timed region: %d k_step launches x %d envs, algorithmic bytes / region time '
                                        '(planner stand-in launches inside that time)' % (K * n_groups, E // n_groups),
                     'achieved_standalone': achieved_alone, 'frac_standalone': achieved_alone / peak,
                     'kernel_ms': ms_step_kernel,
                     'kernel_ms_source': 'CUDA events around each k_step launch over the %d steps that follow the '
                                         'timed region (same episodes; launched stand-alone: in the timed region the '
                                         'kernel starts under the planner stand-in, DCA_OPT_OVERLAP_PLANNER)' % Kfirst,
                     'kernel_ms_episode_avg': ms_step_kernel_episode, 'episode_avg_steps': Kk,
                     'timed_region_ms_per_step': ms_total / K,
                     'frac_timed_region': algo_bytes / (ms_total / K * 1e-3) / 1e9 / peak,
                     'timed_region_note': 'in the timed region launches overlap (env groups on their own streams / '
                                          'launch t+1 starting inside the last wave of launch t), so a launch has no '
                                          'duration of its own there: this is the algorithmic bytes of all k_step '
                                          'launches of the region / region time (the planner stand-in launches are '
                                          'inside that time); kernel_ms and frac are the same kernel launched '
                                          'stand-alone over all %d envs' % E,
                     'pipelined_ms_per_step': one_group['ms_per_step'],
                     'frac_pipelined': algo_bytes / (one_group['ms_per_step'] * 1e-3) / 1e9 / peak,
                     'kernel_ms_synchronised_start': ms_step_sync_start,
                     'frac_synchronised_start': algo_bytes / (ms_step_sync_start * 1e-3) / 1e9 / peak,
                     'frac_episode_avg': algo_bytes / (ms_step_kernel_episode * 1e-3) / 1e9 / peak,
                     'dram_frac': dram_frac(ncu, 'k_step', ms_step_kernel_episode, peak, E, args.scenario),
                     'dram_frac_timed_region': (dram_frac(ncu, 'k_step_two_groups', ms_total / K, peak, E, args.scenario)
                                                if n_groups == 2 else None),
                     'dram_frac_timed_region_source': ncu.get('range_source_two_groups') if n_groups == 2 else None,
                     'traffic_source': ncu.get('source')},
        'roofline_gather_kernel': {'bound': 'hbm', 'kernel': 'k_observe', 'unit': 'GB/s', 'peak': peak,
                                   'achieved': gather_bytes / (ms_gather * 1e-3) / 1e9,
                                   'frac': gather_bytes / (ms_gather * 1e-3) / 1e9 / peak,
                                   'algorithmic_bytes_per_launch': gather_bytes, 'kernel_ms': ms_gather,
                                   'traffic': ncu.get('k_observe_dram_bytes_per_launch'),
                                   'dram_frac': dram_frac(ncu, 'k_observe', ms_gather, peak, E, args.scenario),
                                   'dram_frac_source': ncu.get('range_source'),
                                   'two_groups': {
                                       'ms_per_pass': ms_gather_groups,
                                       'frac': gather_bytes / (ms_gather_groups * 1e-3) / 1e9 / peak,
                                       'dram_frac': dram_frac(ncu, 'k_observe_two_groups', ms_gather_groups, peak, E, args.scenario),
                                       'what': 'the same pass over all %d envs as two groups of %d on their own streams '
                                               '(mid-episode state, step 60 of synchronised first episodes)' % (E, E // 2),
                                       'dram_frac_source': ncu.get('range_source_two_groups')}},
        'histpool_kernel': {'bound': 'hbm', 'kernel': 'k_histpool', 'unit': 'GB/s', 'kernel_ms': ms_hist,
                            'achieved': hist_bytes / (ms_hist * 1e-3) / 1e9, 'peak': peak,
                            'frac': hist_bytes / (ms_hist * 1e-3) / 1e9 / peak},
        'runner_update_ms': ms_runner,
        'multiteam_kernel': {'bound': 'hbm', 'kernel': 'k_mt_step', 'workload': 'dca_multiteam 25v25, terrain on, jam 0.05, %d envs' % E,
                             'kernel_ms': ms_mt, 'agent_steps_per_sec': E * mt_N / (ms_mt * 1e-3), 'unit': 'GB/s',
                             'achieved': mt_bytes / (ms_mt * 1e-3) / 1e9, 'peak': peak,
                             'frac': mt_bytes / (ms_mt * 1e-3) / 1e9 / peak, 'alive_fraction_at_end': mt_alive,
                             'algorithmic_bytes_per_launch': mt_bytes,
                             'two_groups': {'ms_per_step': ms_mt_groups,
                                            'agent_steps_per_sec': E * mt_N / (ms_mt_groups * 1e-3),
                                            'frac': mt_bytes / (ms_mt_groups * 1e-3) / 1e9 / peak,
                                            'what': 'the same envs stepped as two groups of %d on their own streams' % (E // 2)},
                             'traffic': ncu.get('k_mt_step_dram_bytes_per_launch') if E == ncu.get('envs_per_gpu') else None},
        'gae_kernel': {'bound': 'hbm', 'kernel': 'k_gae', 'unit': 'GB/s',
                       'shape_TEG': [T, E, G], 'kernel_ms': ms_gae, 'achieved': gae_bytes / (ms_gae * 1e-3) / 1e9,
                       'peak': peak, 'frac': gae_bytes / (ms_gae * 1e-3) / 1e9 / peak},
        'e2e': {'value': e2e_value, 'unit': 'agent-steps/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'steps': Ke, 'api': 'DcaSuperpoolEnv.step (numpy actions in; planner + step on device; numpy '
                                    'obs/reward/done/info out)'},
        'one_group': one_group if grouped is not None else None,
        'with_scripted_planner': scripted,
        'other_configs': other,
        'policy_rollout_50v55': policy_line,
        'collectives': coll,
        'gpu_launches': gpu_launches,
        'clocks': clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_cpu_env, cpu_steps = 1024, 60
        v, dt = cpu_port_run(n_cpu_env, cpu_steps, 5, cores)
        v0, _ = cpu_port_run(n_cpu_env, cpu_steps, 5, cores, full_planner=False)
        v1, dt1 = cpu_port_run(128, 40, 3, 1)
        ref = reference_pool_run(args.scenario, 360, 5, procs=cores, fold=2)  # two whole episodes, ~10 s
        if ref is not None:  # the reference's own path (SmartPool + SuperpoolEnv + MISSION/dca), staged in oracle/_ref
            line['cpu_baseline'] = {
                'value': ref['agent_steps_per_sec'], 'unit': 'agent-steps/s', 'cores': cores, 'kind': 'reference',
                'sample': '%d envs (%d worker processes x fold %d) x 360 steps (two whole episodes) of the same %s workload through the '
                          'unmodified SuperpoolEnv.step, scripted attacker planner inside World.step (%.1f s)'
                          % (ref['envs'], ref['procs'], ref['fold'], args.scenario, ref['seconds']),
                'port_value': v, 'port_value_without_planner': v0, 'port_single_core_value': v1,
                'port_over_reference': v / ref['agent_steps_per_sec']}
        else:
            line['cpu_baseline'] = {'value': v, 'unit': 'agent-steps/s', 'cores': cores, 'kind': 'port',
                                    'sample': '%d envs x %d steps of the same workload incl. the scripted attacker '
                                              'planner, C/OpenMP port (oracle/_ref not staged) (%.1f s)'
                                              % (n_cpu_env, cpu_steps, dt),
                                    'single_core_value': v1, 'value_without_planner': v0}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    global WORKLOAD
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=int(os.environ.get('WORLD_SIZE', '1')))
    ap.add_argument('--steps', type=int, default=1800)
    ap.add_argument('--warmup', type=int, default=180)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--envs-per-gpu', type=int, default=ENVS_PER_GPU)
    ap.add_argument('--e2e-steps', type=int, default=40)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-other-configs', action='store_true', help='skip the 100v100 / 50v55 report lines')
    ap.add_argument('--planner', default='centroid', choices=['centroid', 'scripted'],
                    help='attacker forces: the centroid stand-in of configs[1] (default) or the scripted formation planner k_plan')
    ap.add_argument('--env-groups', type=int, default=0,
                    help='step the envs as this many independent groups on their own streams (DcaEnvGroups); '
                         '0 = 2 with the centroid stand-in at 4096 envs per GPU or more, else 1')
    ap.add_argument('--no-preroll', action='store_true',
                    help='time from a synchronised reset of all envs instead of the stationary episode-phase mix')
    ap.add_argument('--no-pipeline', action='store_true',
                    help='DCA_OPT_PIPELINE_STEPS off: every launch waits for the whole previous launch')
    ap.add_argument('--no-overlap', action='store_true',
                    help='launch the step kernel strictly after the planner stand-in (DCA_OPT_OVERLAP_PLANNER off)')
    ap.add_argument('--traffic', type=float, default=None,
                    help='dram bytes per k_step launch from the committed ncu --set full capture (profiles/)')
    ap.add_argument('--scenario', default='50v50', choices=sorted(SCENARIOS),
                    help='50v50 = BASELINE.json configs[1] (the headline); the others are extra report lines')
    args = ap.parse_args()
    WORKLOAD = dict(WORKLOAD, **SCENARIOS[args.scenario])
    args.warmup = max(args.warmup, 3)
    if args.impl == 'reference':
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
