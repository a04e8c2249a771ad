"""Vectorised DCA env on one B200: the drop-in for the reference's SuperpoolEnv / SmartPool / EnvAutoReset stack.

Two facades over the C-ABI (include/dca_b200.h):

* ``DcaVecEnv``       -- "fast": everything stays on the device; step() returns the bound torch CUDA tensors
                         (zero copy) for a torch policy / rollout loop.
* ``DcaSuperpoolEnv`` -- "compat": the exact host surface of /root/reference/UTIL/shm_env.py:126-168
                         (numpy in, numpy + per-env info dicts out, NaN action row = ENV-PAUSE), so unmodified
                         ALGORITHM modules (conc, conc_4hist, random) and task_runner.Runner can drive it.
``make_parallel_envs`` mirrors /root/reference/MISSION/env_router.py:68-117.

There is no CPU path: constructing either class without a CUDA device / without libdca_b200.so raises.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .config import GlobalConfig, ScenarioConfig

OBS_ROWS, OBS_ROW = 12, 16


def pod_config(scfg=ScenarioConfig, seed=0, env_gid0=0):
    """POD snapshot of ScenarioConfig, taken at construction time (parallel_run.py:30-101)."""
    tp = list(scfg.terrain_parameters)
    return _lib.DcaConfig(
        int(scfg.num_guards), int(scfg.num_attackers), float(scfg.agent_acc), float(scfg.agent_max_speed),
        float(scfg.agent_max_rot), float(scfg.agent_shoot_base_radius), float(scfg.agent_shoot_win),
        float(scfg.random_jam_prob), int(bool(scfg.introduce_terrain)), float(tp[0]), float(tp[1]),
        int(scfg.MaxEpisodeStep), float(scfg.init_distance), int(bool(scfg.enable_attacker_fluctuation)),
        int(scfg.attacker_fluctuation), float(scfg.size), int(bool(scfg.half_death_reward)),
        int(bool(scfg.REWARD_DEBUG)), float(scfg.REWARD_DEBUG_Value), int(bool(scfg.DISALBE_RED_FUNCTION)),
        int(seed) & 0xFFFFFFFFFFFFFFFF, int(env_gid0))


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class DcaVecEnv(object):
    """n_env DCA envs resident on one GPU.  All tensors returned are views of buffers written in place."""

    def __init__(self, n_env, scfg=ScenarioConfig, device=None, seed=0, env_gid0=0, debug_outputs=False,
                 empty_rows_nan=False, overlap_planner=True, pipeline_steps=False, buffers=None):
        if not torch.cuda.is_available():
            raise _lib.DcaError('hmp2g_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback')
        self.lib = _lib.load()
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        assert self.device.type == 'cuda'
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device('cuda', dev_index)
        self.cfg = pod_config(scfg, seed, env_gid0)
        self.E, self.G, self.A = int(n_env), self.cfg.num_guards, self.cfg.num_attackers
        self.N = self.G + self.A
        self.num_envs = self.E
        self.handle = C.c_void_p()
        _lib.check(self.lib.dca_create(C.byref(self.cfg), self.E, dev_index, C.byref(self.handle)))
        E, G, A, N, dev = self.E, self.G, self.A, self.N, self.device
        def out(name, shape, dtype):  # `buffers`: caller-owned output tensors (DcaEnvGroups binds slices of its own)
            if buffers is None or name not in buffers:
                return torch.zeros(shape, dtype=dtype, device=dev)
            self._check_tensor(buffers[name], dtype, shape, name)
            return buffers[name]
        self.obs = out('obs', (E, G, OBS_ROWS, OBS_ROW), torch.float32)
        self.tobs = out('tobs', (E, G, OBS_ROWS, OBS_ROW), torch.float32)
        self.reward = out('reward', (E, G), torch.float32)
        self.done = out('done', (E,), torch.uint8)
        self.win = out('win', (E,), torch.uint8)
        self.alive = out('alive', (E, N), torch.uint8)
        self.slot_id = torch.full((E, G, OBS_ROWS), -1, dtype=torch.int16, device=dev) if debug_outputs else None
        self.nearest = torch.full((E, A), -1, dtype=torch.int32, device=dev) if debug_outputs else None
        bufs = _lib.DcaBuffers(_ptr(self.obs), _ptr(self.tobs), _ptr(self.reward), _ptr(self.done), _ptr(self.win),
                               _ptr(self.alive), _ptr(self.slot_id), _ptr(self.nearest))
        _lib.check(self.lib.dca_bind_buffers(self.handle, C.byref(bufs)), self.handle)
        self._force = torch.zeros((E, A, 2), dtype=torch.float64, device=dev)
        if empty_rows_nan:  # hand observations over the way conc's ShellEnvWrapper makes them (shell_env.py:72-75)
            _lib.check(self.lib.dca_set_option(self.handle, 1, 1), self.handle)
        # DCA_OPT_OVERLAP_PLANNER: a step issued right behind plan() / plan_centroid() starts under the planner kernel
        _lib.check(self.lib.dca_set_option(self.handle, 2, int(bool(overlap_planner))), self.handle)
        # DCA_OPT_PIPELINE_STEPS: plan / step / plan / step ... issued back to back hand envs over one by one
        _lib.check(self.lib.dca_set_option(self.handle, 4, int(bool(pipeline_steps))), self.handle)

    # -- lifetime -----------------------------------------------------------------------
    def close(self):
        if getattr(self, 'handle', None) is not None and self.handle.value:
            self.lib.dca_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _check_tensor(self, t, dtype, shape, name):
        assert t.is_cuda and t.device == self.device, '%s must live on %s' % (name, self.device)
        assert t.dtype == dtype and tuple(t.shape) == tuple(shape) and t.is_contiguous(), \
            '%s must be contiguous %s%s, got %s%s' % (name, dtype, tuple(shape), t.dtype, tuple(t.shape))

    # -- env surface --------------------------------------------------------------------
    def reset(self, env_mask=None):
        """collective_assaultGlobalEnv.reset (parallel_run.py:248-270) for all envs (or the masked ones)."""
        if env_mask is not None:
            self._check_tensor(env_mask, torch.uint8, (self.E,), 'env_mask')
        _lib.check(self.lib.dca_reset(self.handle, _ptr(env_mask), self._stream()), self.handle)
        return self.obs

    def plan_centroid(self, out=None):
        """Benchmark planner stand-in: fills (and returns) the attacker-force tensor f64 [E,A,2]."""
        out = self._force if out is None else out
        self._check_tensor(out, torch.float64, (self.E, self.A, 2), 'att_force')
        _lib.check(self.lib.dca_plan_centroid(self.handle, _ptr(out), self._stream()), self.handle)
        return out

    def plan(self, out=None, team_label=None, team_centroid=None):
        """The scripted attacker formation planner (cheat_script_ai.py:19-194) on the resident state: fills (and
        returns) the attacker-force tensor f64 [E,A,2].  team_label int32 [E,A] + team_centroid f64 [E,A//5,2] inject
        a k-means result (parity harness); by default the teams come from the keyed k-means++ in the kernel."""
        out = self._force if out is None else out
        self._check_tensor(out, torch.float64, (self.E, self.A, 2), 'att_force')
        if team_label is not None:
            K = max(self.A // 5, 1)
            self._check_tensor(team_label, torch.int32, (self.E, self.A), 'team_label')
            self._check_tensor(team_centroid, torch.float64, (self.E, K, 2), 'team_centroid')
        _lib.check(self.lib.dca_plan(self.handle, _ptr(out), _ptr(team_label), _ptr(team_centroid), self._stream()),
                   self.handle)
        return out

    def step(self, actions, att_force=None, skip=None, rng_words=None, autoreset=True):
        """actions int32 [E,G]; att_force float64 [E,A,2] (None -> the scripted attacker planner, plan());
        skip uint8 [E]; rng_words int32/uint32-as-int32 [E,G,2,12].  Returns (obs, reward, done, win) views."""
        self._check_tensor(actions, torch.int32, (self.E, self.G), 'actions')
        if att_force is None:
            att_force = self.plan()
        else:
            self._check_tensor(att_force, torch.float64, (self.E, self.A, 2), 'att_force')
        if skip is not None:
            self._check_tensor(skip, torch.uint8, (self.E,), 'skip')
        if rng_words is not None:
            self._check_tensor(rng_words, torch.int32, (self.E, self.G, 2, 12), 'rng_words')
        _lib.check(self.lib.dca_step(self.handle, _ptr(actions), _ptr(att_force), _ptr(skip), _ptr(rng_words),
                                     int(bool(autoreset)), self._stream()), self.handle)
        return self.obs, self.reward, self.done, self.win

    def observe(self, rng_words=None):
        if rng_words is not None:
            self._check_tensor(rng_words, torch.int32, (self.E, self.G, 2, 12), 'rng_words')
        _lib.check(self.lib.dca_observe(self.handle, _ptr(rng_words), self._stream()), self.handle)
        return self.obs

    def kernel_launches(self):
        n = C.c_int64()
        _lib.check(self.lib.dca_kernel_launches(self.handle, C.byref(n)), self.handle)
        return n.value

    # -- state import / export (parity harness, checkpoints) ----------------------------------
    _STATE_SPEC = (('pos', torch.float64, 2), ('vel', torch.float64, 2), ('atk_rad', torch.float64, 0),
                   ('terrain', torch.float64, 0), ('alive', torch.uint8, 0), ('just_died', torch.uint8, 0),
                   ('hit', torch.uint8, 0), ('was_hit', torch.uint8, 0), ('num_hit', torch.int32, 0),
                   ('num_was_hit', torch.int32, 0))
    _ENV_SPEC = (('init_theta', torch.float64), ('time_step', torch.int32), ('n_alive_guards', torch.int32),
                 ('n_alive_attackers', torch.int32), ('obs_counter', torch.int32), ('episode', torch.int32))

    def _alloc_state(self):
        st = {}
        for name, dt, k in self._STATE_SPEC:
            st[name] = torch.zeros((self.E, self.N, k) if k else (self.E, self.N), dtype=dt, device=self.device)
        for name, dt in self._ENV_SPEC:
            st[name] = torch.zeros((self.E,), dtype=dt, device=self.device)
        return st

    def export_state(self):
        """dict of numpy arrays in a plain SoA layout (counters as uint32)."""
        st = self._alloc_state()
        s = _lib.DcaState(*[_ptr(st[n]) for n in _lib.STATE_FIELDS])
        _lib.check(self.lib.dca_export_state(self.handle, C.byref(s), self._stream()), self.handle)
        torch.cuda.current_stream(self.device).synchronize()
        out = {k: v.cpu().numpy() for k, v in st.items()}
        out['obs_counter'] = out['obs_counter'].view(np.uint32)
        out['episode'] = out['episode'].view(np.uint32)
        return out

    def import_state(self, state):
        """state: dict of numpy arrays [E,...] (same keys as export_state)."""
        st = self._alloc_state()
        for k, t in st.items():
            a = np.asarray(state[k])
            if k in ('obs_counter', 'episode'):
                a = a.astype(np.int64).astype(np.uint32).view(np.int32)
            np_dt = {torch.float64: np.float64, torch.uint8: np.uint8, torch.int32: np.int32}[t.dtype]
            t.copy_(torch.from_numpy(np.ascontiguousarray(a).astype(np_dt)).reshape(t.shape))
        s = _lib.DcaState(*[_ptr(st[n]) for n in _lib.STATE_FIELDS])
        _lib.check(self.lib.dca_import_state(self.handle, C.byref(s), self._stream()), self.handle)
        torch.cuda.current_stream(self.device).synchronize()


class DcaEnvGroups(object):
    """n_env envs as `groups` independent DcaVecEnv of n_env / groups envs each -- contiguous global env ids, so the
    results are those of one DcaVecEnv(n_env) bit for bit (every draw is keyed by the global env id) -- issuing on their
    own CUDA streams and writing their slices of shared [n_env, ...] output tensors.

    Why: a launch of E envs ends in a tail (the last, partial wave of CTAs drains while nothing else may start); the
    launch chains of the groups are independent, so one group's CTAs fill the slots the other group's launch boundary
    leaves empty, and a latency-bound kernel of one group (the scripted planner) runs under the other group's work.
    The classic double-buffered sampler: with a policy in the loop, group A's policy forward overlaps group B's env step.

    step(..., sync=True) orders the launches behind the caller's current stream (fork) and the caller's stream behind
    the results (join).  sync=False only launches: the caller brackets a burst of steps with fork() / join() itself
    (actions that do not depend on the observations, e.g. a pre-generated pool)."""

    def __init__(self, n_env, groups=2, env_gid0=0, device=None, **kw):
        assert n_env % groups == 0, 'n_env must be a multiple of the group count'
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device('cuda', torch.cuda.current_device())
        self.E, self.groups = int(n_env), int(groups)
        self.Eg = self.E // self.groups
        self.num_envs = self.E
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(self.groups)]
        self.envs = []
        first = DcaVecEnv(self.Eg, env_gid0=env_gid0, device=self.device, **kw)  # its own buffers: gives the shapes
        self.G, self.A, self.N = first.G, first.A, first.N
        full = lambda t: torch.zeros((self.E,) + tuple(t.shape[1:]), dtype=t.dtype, device=self.device)
        self.obs, self.tobs, self.reward = full(first.obs), full(first.tobs), full(first.reward)
        self.done, self.win, self.alive = full(first.done), full(first.win), full(first.alive)
        first.close()
        torch.cuda.synchronize(self.device)
        for k in range(self.groups):
            sl = slice(k * self.Eg, (k + 1) * self.Eg)
            bufs = dict(obs=self.obs[sl], tobs=self.tobs[sl], reward=self.reward[sl], done=self.done[sl],
                        win=self.win[sl], alive=self.alive[sl])
            self.envs.append(DcaVecEnv(self.Eg, env_gid0=env_gid0 + k * self.Eg, device=self.device, buffers=bufs, **kw))
        self._ev = [torch.cuda.Event() for _ in range(self.groups)]

    def close(self):
        for e in self.envs:
            e.close()

    def fork(self):
        """Group streams wait for everything issued so far on the caller's current stream."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        for s in self.streams:
            s.wait_event(ev)

    def join(self):
        """The caller's current stream waits for everything the groups have issued."""
        cur = torch.cuda.current_stream(self.device)
        for k, s in enumerate(self.streams):
            self._ev[k].record(s)
            cur.wait_event(self._ev[k])

    def _each(self, fn, sync):
        if sync:
            self.fork()
        for k, env in enumerate(self.envs):
            with torch.cuda.stream(self.streams[k]):
                fn(k, env, slice(k * self.Eg, (k + 1) * self.Eg))
        if sync:
            self.join()

    def reset(self, sync=True):
        self._each(lambda k, env, sl: env.reset(), sync)
        return self.obs

    def step(self, actions, att_force=None, planner='scripted', skip=None, autoreset=True, sync=True):
        """actions int32 [E,G]; att_force f64 [E,A,2] or None -> planner 'scripted' (k_plan) / 'centroid' (stand-in) per
        group on its stream.  Returns (obs, reward, done, win): the shared tensors."""
        def one(k, env, sl):
            f = att_force[sl] if att_force is not None else (env.plan() if planner == 'scripted' else env.plan_centroid())
            env.step(actions[sl], f, skip=None if skip is None else skip[sl], autoreset=autoreset)
        self._each(one, sync)
        return self.obs, self.reward, self.done, self.win

    def observe(self, sync=True):
        """The observation pass alone on the resident state (DcaVecEnv.observe) for every group."""
        self._each(lambda k, env, sl: env.observe(), sync)
        return self.obs

    def kernel_launches(self):
        return sum(e.kernel_launches() for e in self.envs)


class DcaSuperpoolEnv(object):
    """Host-facing drop-in for UTIL.shm_env.SuperpoolEnv (UTIL/shm_env.py:126-168) backed by DcaVecEnv.

    step(actions float64 [E,G], NaN row => skip) -> (obs f32 [E,G,192], reward f64 [E,G], done bool [E],
    info object[E] of dict).  info = {} | {'win': bool}; on done also 'obs-echo' (terminal obs [G,192]) and
    'others' (reset info).  The returned obs of a done env is the post-reset obs (shm_env.py:51-66).
    A skipped env returns its previous (ob, reward, done, info) (shm_env.py:37-41).
    Attacker forces come from the scripted planner kernel (DcaVecEnv.plan, cheat_script_ai.py:19-194), as in the
    reference where World.step runs it (core.py:189-196), or from `planner` (callable(env) -> f64 cuda [E,A,2]).
    """

    def __init__(self, process_pool=None, env_args_dict_list=None, spaces=None, n_env=None, scfg=None,
                 device=None, seed=None, planner=None, env_gid0=0, reference_obs_dtype=False):
        scfg = scfg or GlobalConfig.ScenarioConfig
        if n_env is None:
            n_env = len(env_args_dict_list) if env_args_dict_list is not None else GlobalConfig.num_threads
        self.SuperPool = process_pool  # unused: there are no worker processes
        self.num_envs = n_env
        self.env_name_marker = env_args_dict_list[0][0]['marker'] if env_args_dict_list else ''
        seed = GlobalConfig.seed if seed is None else seed
        self.env = DcaVecEnv(n_env, scfg=scfg, device=device, seed=seed, env_gid0=env_gid0)
        self.planner = planner
        # reference_obs_dtype: hand observations back as float64 once any guard of the batch is dead, as the reference
        # does (a dead guard's observation is np.zeros(192), float64, and np.stack promotes the batch);
        # the values are float32 numbers either way.  Default: always float32 (no 2x host copy).
        self.reference_obs_dtype = bool(reference_obs_dtype)
        self.observation_space, self.action_space = [], []  # parallel_run.py:170-171
        E, G = self.env.E, self.env.G
        pin = dict(pin_memory=True)
        self._h_act = torch.zeros((E, G), dtype=torch.int32, **pin)
        self._h_skip = torch.zeros((E,), dtype=torch.uint8, **pin)
        self._d_act = torch.zeros((E, G), dtype=torch.int32, device=self.env.device)
        self._d_skip = torch.zeros((E,), dtype=torch.uint8, device=self.env.device)
        # two pinned landing buffers, used alternately: the arrays handed out by step t stay valid while step t+1
        # runs (the reference hands out fresh arrays; callers copy what they keep, task_runner.py:100)
        self._h_obs_pair = [torch.zeros((E, G, OBS_ROWS * OBS_ROW), dtype=torch.float32, **pin) for _ in range(2)]
        self._flip = 0
        self._h_obs = self._h_obs_pair[0]
        self._h_tobs = torch.zeros((E, G, OBS_ROWS * OBS_ROW), dtype=torch.float32, **pin)
        self._h_rew = torch.zeros((E, G), dtype=torch.float32, **pin)
        self._h_done = torch.zeros((E,), dtype=torch.uint8, **pin)
        self._h_win = torch.zeros((E,), dtype=torch.uint8, **pin)
        self._info = np.empty((E,), dtype=object)
        for i in range(E):
            self._info[i] = {}
        self._info_dirty = np.zeros((E,), dtype=bool)
        self._cold_start = True
        self._suffer_reset = np.zeros((E,), dtype=bool)
        # the planner reads only the resident state, so the plan for step t+1 is computed right after step t on a
        # second stream, under the device->host copy of step t's observations
        self._plan_stream = torch.cuda.Stream(device=self.env.device)
        self._plan_ready = None
        self.h2d_bytes = self._h_act.numel() * 4 + self._h_skip.numel()
        self.d2h_bytes = (self._h_obs.numel() + self._h_rew.numel()) * 4 + 2 * E

    def get_space(self):
        return {'obs_space': self.observation_space, 'act_space': self.action_space}

    def _plan_ahead(self):
        if self.planner is not None:
            return
        env = self.env
        self._plan_stream.wait_stream(torch.cuda.current_stream(env.device))
        with torch.cuda.stream(self._plan_stream):
            env.plan()
            self._plan_ready = torch.cuda.Event()
            self._plan_ready.record(self._plan_stream)

    def _take_plan(self):
        """The force tensor planned ahead for the current state (None: nothing valid is in flight)."""
        if self._plan_ready is None:
            return None
        torch.cuda.current_stream(self.env.device).wait_event(self._plan_ready)
        self._plan_ready = None
        return self.env._force

    def _fetch(self):
        env = self.env
        self._flip ^= 1
        self._h_obs = self._h_obs_pair[self._flip]
        self._h_obs.copy_(env.obs.view(env.E, env.G, -1), non_blocking=True)
        self._h_rew.copy_(env.reward, non_blocking=True)
        self._h_done.copy_(env.done, non_blocking=True)
        self._h_win.copy_(env.win, non_blocking=True)
        torch.cuda.current_stream(env.device).synchronize()

    def reset(self):
        if self._cold_start:  # EnvAutoReset.reset shm_env.py:82-93
            self._cold_start = False
            self._take_plan()
            self.env.reset()
            self._plan_ahead()
        elif not self._suffer_reset.all():
            self._take_plan()  # the state changes under it: wait for the plan in flight, then drop it
            mask = torch.from_numpy((~self._suffer_reset).astype(np.uint8)).to(self.env.device)
            self.env.reset(mask)
            self._plan_ahead()
        self._suffer_reset[:] = True
        self._fetch()
        info = np.empty((self.num_envs,), dtype=object)
        for i in range(self.num_envs):
            info[i] = {'others': None}
        return self._h_obs.numpy(), info

    def step(self, actions):
        env = self.env
        actions = np.asarray(actions)
        if actions.ndim == 3 and actions.shape[-1] == 1:  # RandomController layout (ALGORITHM/random/foundation.py:28)
            actions = actions[..., 0]
        skip = np.isnan(actions.sum(axis=1))  # a NaN anywhere in the row (ENV-PAUSE, shm_env.py:37)
        if skip.any():
            assert not (skip & ~self._suffer_reset).any(), 'skip command on an env that was not just reset'
            actions = np.where(np.isnan(actions), 0, actions)
        self._h_skip.numpy()[:] = skip
        np.copyto(self._h_act.numpy(), actions, casting='unsafe')  # float64 holding ints -> int32, into pinned memory
        self._d_act.copy_(self._h_act, non_blocking=True)
        self._d_skip.copy_(self._h_skip, non_blocking=True)
        force = self.planner(env) if self.planner is not None else self._take_plan()
        env.step(self._d_act, force, skip=self._d_skip)
        self._plan_ahead()
        self._fetch()
        done = self._h_done.numpy().astype(bool)
        win = self._h_win.numpy().astype(bool)
        stepped = ~skip
        info = self._info
        idx_done = np.nonzero(stepped & done)[0]
        for i in np.nonzero(stepped & ~done & self._info_dirty)[0]:
            info[i] = {}
        self._info_dirty[stepped] = done[stepped]
        if len(idx_done):
            sel = torch.from_numpy(idx_done).to(env.device)
            tobs = env.tobs.view(env.E, env.G, -1)[sel].cpu().numpy()
            for k, i in enumerate(idx_done):
                info[i] = {'win': bool(win[i]), 'obs-echo': tobs[k], 'others': None}
        self._suffer_reset[stepped] = done[stepped]
        obs = self._h_obs.numpy()
        if self.reference_obs_dtype and (obs[:, :, 0] == 0).any():  # a guard whose own record is empty is dead
            obs = obs.astype(np.float64)
        return (obs, self._h_rew.numpy().astype(np.float64), done, info.copy())

    def sleep(self):
        return None


def make_parallel_envs(process_pool=None, marker=''):
    """MISSION.env_router.make_parallel_envs (env_router.py:68-117) for env_name 'dca' and 'dca_multiteam'."""
    assert GlobalConfig.env_name in ('dca', 'collective_assault', 'dca_multiteam'), GlobalConfig.env_name
    env_args_dict_list = [({'env_name': GlobalConfig.env_name,
                            'proc_index': i if 'test' not in marker else -(i + 1),
                            'marker': marker},) for i in range(GlobalConfig.num_threads)]
    if GlobalConfig.env_name == 'dca_multiteam':
        from .multiteam import DcaMultiTeamSuperpoolEnv, MtScenarioConfig
        return DcaMultiTeamSuperpoolEnv(process_pool, env_args_dict_list, scfg=MtScenarioConfig, seed=GlobalConfig.seed)
    return DcaSuperpoolEnv(process_pool, env_args_dict_list)
