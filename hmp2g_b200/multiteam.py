"""dca_multiteam on the GPU: host mirror of /root/reference/MISSION/dca_multiteam.

MtScenarioConfig mirrors the reference's ScenarioConfig (collective_assault_parallel_run.py:27-77) field by field;
DcaMultiTeamVecEnv is the device façade (torch tensors written in place by k_mt_step), DcaMultiTeamSuperpoolEnv the
host façade with the SuperpoolEnv / EnvAutoReset surface (UTIL/shm_env.py:33-97, :126-168): numpy float64 actions for
every agent in (NaN row = ENV-PAUSE), numpy obs / reward / done / info out, info = {} or {'team_ranking', 'end_reason',
'obs-echo'} on done.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .config import ChainVar


def arrange_id(n_agent_each_team):
    """UTIL/tensor_ops.arrange_id: contiguous uid ranges per team."""
    out, begin = [], 0
    for n in n_agent_each_team:
        out.append(range(begin, begin + n))
        begin += n
    return out


class MtScenarioConfig(object):
    """MISSION/dca_multiteam/collective_assault_parallel_run.py:27-77 (same names, same defaults)."""
    N_TEAM = 2
    N_AGENT_EACH_TEAM = [25, 25]
    AGENT_ID_EACH_TEAM = arrange_id(N_AGENT_EACH_TEAM)
    AGENT_ID_EACH_TEAM_CV = ChainVar(lambda N_AGENT_EACH_TEAM: arrange_id(N_AGENT_EACH_TEAM),
                                     chained_with=['N_AGENT_EACH_TEAM'])
    TEAM_NAMES = ['ALGORITHM.Starcraft.star_foundation->StarFoundation',
                  'ALGORITHM.Starcraft.star_foundation->StarFoundation']
    size = 5.0
    random_jam_prob = 0.0
    introduce_terrain = False
    terrain_parameters = [0, 0]
    MaxEpisodeStep = 150
    StepsToNarrowCircle = 100
    init_distance = 2.5
    render = False
    render_with_unity = False
    map_ = 'default'
    benchmark = False
    obs_vec_length = 17
    StateProvided = False
    AvailActProvided = False
    RewardAsUnity = False
    ObsAsUnity = False
    EntityOriented = True
    MCOM_DEBUG = False
    Terrain_DEBUG = False
    half_death_reward = True
    n_actions = 7


def pod_mt_config(scfg, seed=0, env_gid0=0):
    sizes = list(scfg.N_AGENT_EACH_TEAM)
    c = _lib.DcaMtConfig()
    c.n_team = len(sizes)
    if c.n_team > 8:
        raise _lib.DcaError('at most 8 teams')
    for i, n in enumerate(sizes):
        c.team_size[i] = int(n)
    c.size, c.random_jam_prob = float(scfg.size), float(scfg.random_jam_prob)
    c.introduce_terrain = int(bool(scfg.introduce_terrain))
    c.terrain_A, c.terrain_B = float(scfg.terrain_parameters[0]), float(scfg.terrain_parameters[1])
    c.max_episode_step, c.steps_to_narrow = int(scfg.MaxEpisodeStep), int(scfg.StepsToNarrowCircle)
    c.init_distance, c.half_death_reward = float(scfg.init_distance), int(bool(scfg.half_death_reward))
    c.seed, c.env_gid0 = int(seed), int(env_gid0)
    return c


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class DcaMultiTeamVecEnv(object):
    OBS_ROWS, OBS_ROW = 12, 17

    def __init__(self, n_env, scfg=None, device=None, seed=0, env_gid0=0, debug_outputs=False):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.DcaError('hmp2g_b200 needs a CUDA device; there is no CPU fallback')
        self.device = torch.device(device if device is not None else 'cuda')
        assert self.device.type == 'cuda'
        if self.device.index is None:  # 'cuda' means the current device, exactly as DcaVecEnv resolves it
            self.device = torch.device('cuda', torch.cuda.current_device())
        scfg = scfg or MtScenarioConfig
        self.cfg = pod_mt_config(scfg, seed, env_gid0)
        self.E, self.n_team = int(n_env), self.cfg.n_team
        self.team_size = [self.cfg.team_size[i] for i in range(self.n_team)]
        self.N = sum(self.team_size)
        self.agent_id_each_team = arrange_id(self.team_size)
        # team-level interface (collective_assault_parallel_run.py:166-186): RewardAsUnity -> one reward per team, 0
        # (sparse_reward, envs/collective_assault_env.py:77-78) plus the ranking bonus on done (_get_done :245-266);
        # ObsAsUnity -> only agent 0's observation is handed out.  The kernel computes the per-agent quantities either
        # way (the keyed observation draws of agent 0 do not depend on the others being observed).
        self.reward_as_unity = bool(getattr(scfg, 'RewardAsUnity', False))
        self.obs_as_unity = bool(getattr(scfg, 'ObsAsUnity', False))
        self.handle = C.c_void_p()
        _lib.check(self.lib.dca_mt_create(C.byref(self.cfg), self.E, self.device.index, C.byref(self.handle)))
        E, N, dev = self.E, self.N, self.device
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        self.obs = z((E, N, 12, 17), torch.float32)
        self.tobs = z((E, N, 12, 17), torch.float32)
        self.reward = z((E, N), torch.float32)
        self.done = z((E,), torch.uint8)
        self.ranking = z((E, self.n_team), torch.int32)
        self.alive = z((E, N), torch.uint8)
        self.slot_id = z((E, N, 12), torch.int16) if debug_outputs else None
        self._bufs = _lib.DcaMtBuffers(_ptr(self.obs), _ptr(self.tobs), _ptr(self.reward), _ptr(self.done),
                                       _ptr(self.ranking), _ptr(self.alive), _ptr(self.slot_id))
        _lib.check(self.lib.dca_mt_bind_buffers(self.handle, C.byref(self._bufs)))

    def __del__(self):
        h, self.handle = getattr(self, 'handle', None), None
        if h:
            try:
                self.lib.dca_mt_destroy(h)
            except Exception:
                pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _check(self, t, dt, shape, name):
        if t.dtype != dt or tuple(t.shape) != tuple(shape) or t.device != self.device or not t.is_contiguous():
            raise _lib.DcaError('%s: expected contiguous %s %s on %s, got %s %s on %s'
                                % (name, dt, tuple(shape), self.device, t.dtype, tuple(t.shape), t.device))

    def reset(self, env_mask=None):
        if env_mask is not None:
            self._check(env_mask, torch.uint8, (self.E,), 'env_mask')
        _lib.check(self.lib.dca_mt_reset(self.handle, _ptr(env_mask), self._stream()))
        return self.obs_out()

    def obs_out(self):
        """What the env hands out as observation: every agent's, or agent 0's alone under ObsAsUnity."""
        return self.obs[:, :1] if self.obs_as_unity else self.obs

    def tobs_out(self):
        return self.tobs[:, :1] if self.obs_as_unity else self.tobs

    def reward_out(self):
        """Per-agent reward f32 [E,N], or under RewardAsUnity f32 [E,n_team]: (n_team - 1 - rank) * 10 for a ranked team,
        -10 when every team draws (rank -1), on the step that ends the episode; 0 otherwise."""
        if not self.reward_as_unity:
            return self.reward
        rank = self.ranking.to(torch.float32)
        bonus = torch.where(rank >= 0, (self.n_team - 1 - rank) * 10.0, torch.full_like(rank, -10.0))
        return torch.where((self.done != 0)[:, None], bonus, torch.zeros_like(bonus))

    def step(self, actions, skip=None, autoreset=True):
        """actions int32 [E,N] (one of 0..6 per agent, all teams); skip uint8 [E].
        Returns (obs f32 [E,N,12,17], reward f32 [E,N], done u8 [E], ranking i32 [E,n_team]) -- views, written in place
        (ObsAsUnity: obs [E,1,12,17]; RewardAsUnity: reward [E,n_team], a fresh tensor)."""
        self._check(actions, torch.int32, (self.E, self.N), 'actions')
        if skip is not None:
            self._check(skip, torch.uint8, (self.E,), 'skip')
        _lib.check(self.lib.dca_mt_step(self.handle, _ptr(actions), _ptr(skip), int(bool(autoreset)), self._stream()))
        return self.obs_out(), self.reward_out(), self.done, self.ranking

    _SPEC = (('pos', torch.float64, 2), ('vel', torch.float64, 2), ('atk_rad', torch.float64, 0),
             ('terrain', torch.float64, 0), ('alive', torch.uint8, 0), ('just_died', torch.uint8, 0),
             ('hit', torch.uint8, 0), ('was_hit', torch.uint8, 0), ('bsa', torch.uint8, 0))
    _ENV_SPEC = (('init_theta', torch.float64), ('time_step', torch.int32), ('obs_counter', torch.int32),
                 ('episode', torch.int32))

    def _state_tensors(self):
        t = {n: torch.zeros((self.E, self.N) + ((d,) if d else ()), dtype=dt, device=self.device)
             for n, dt, d in self._SPEC}
        t.update({n: torch.zeros((self.E,), dtype=dt, device=self.device) for n, dt in self._ENV_SPEC})
        return t

    def export_state(self):
        t = self._state_tensors()
        st = _lib.DcaMtState(*[_ptr(t[n]) for n in _lib.MT_STATE_FIELDS])
        _lib.check(self.lib.dca_mt_export_state(self.handle, C.byref(st), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()
        return {k: v.cpu().numpy() for k, v in t.items()}

    def import_state(self, d):
        np_of = {torch.float64: np.float64, torch.uint8: np.uint8, torch.int32: np.int32}
        t = self._state_tensors()
        for k in t:
            a = np.atleast_1d(np.asarray(d[k]))
            if a.dtype == np.uint32:
                a = a.view(np.int32)
            t[k].copy_(torch.from_numpy(np.ascontiguousarray(a.astype(np_of[t[k].dtype])).reshape(tuple(t[k].shape))))
        st = _lib.DcaMtState(*[_ptr(t[n]) for n in _lib.MT_STATE_FIELDS])
        _lib.check(self.lib.dca_mt_import_state(self.handle, C.byref(st), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()


class DcaMultiTeamSuperpoolEnv(object):
    """Host-facing drop-in for UTIL.shm_env.SuperpoolEnv over dca_multiteam envs."""

    def __init__(self, process_pool=None, env_args_dict_list=None, n_env=None, scfg=None, device=None, seed=0,
                 env_gid0=0):
        if n_env is None:
            n_env = len(env_args_dict_list)
        self.num_envs = n_env
        self.env = DcaMultiTeamVecEnv(n_env, scfg=scfg, device=device, seed=seed, env_gid0=env_gid0)
        self.observation_space, self.action_space = [], []
        E, N = self.env.E, self.env.N
        pin = dict(pin_memory=True)
        self._h_act = torch.zeros((E, N), dtype=torch.int32, **pin)
        self._h_skip = torch.zeros((E,), dtype=torch.uint8, **pin)
        self._d_act = torch.zeros((E, N), dtype=torch.int32, device=self.env.device)
        self._d_skip = torch.zeros((E,), dtype=torch.uint8, device=self.env.device)
        n_obs = 1 if self.env.obs_as_unity else N
        n_rew = self.env.n_team if self.env.reward_as_unity else N
        self._h_obs_pair = [torch.zeros((E, n_obs, 204), dtype=torch.float32, **pin) for _ in range(2)]
        self._flip = 0
        self._h_rew = torch.zeros((E, n_rew), dtype=torch.float32, **pin)
        self._h_done = torch.zeros((E,), dtype=torch.uint8, **pin)
        self._h_rank = torch.zeros((E, self.env.n_team), dtype=torch.int32, **pin)
        self._info = np.empty((E,), dtype=object)
        for i in range(E):
            self._info[i] = {}
        self._info_dirty = np.zeros((E,), dtype=bool)
        self._cold_start = True
        self._suffer_reset = np.zeros((E,), dtype=bool)

    def get_space(self):
        return {'obs_space': self.observation_space, 'act_space': self.action_space}

    def _fetch(self):
        env = self.env
        self._flip ^= 1
        self._h_obs = self._h_obs_pair[self._flip]
        self._h_obs.copy_(env.obs_out().reshape(env.E, -1, 204), non_blocking=True)
        self._h_rew.copy_(env.reward_out(), non_blocking=True)
        self._h_done.copy_(env.done, non_blocking=True)
        self._h_rank.copy_(env.ranking, non_blocking=True)
        torch.cuda.current_stream(env.device).synchronize()

    def reset(self):
        if self._cold_start:
            self._cold_start = False
            self.env.reset()
        elif not self._suffer_reset.all():
            self.env.reset(torch.from_numpy((~self._suffer_reset).astype(np.uint8)).to(self.env.device))
        self._suffer_reset[:] = True
        self._fetch()
        info = np.empty((self.num_envs,), dtype=object)
        for i in range(self.num_envs):
            info[i] = {}
        return self._h_obs.numpy(), info

    def step(self, actions):
        env = self.env
        actions = np.asarray(actions, dtype=np.float64)
        skip = np.isnan(actions.sum(axis=1))
        if skip.any():
            assert not (skip & ~self._suffer_reset).any(), 'skip command on an env that was not just reset'
            actions = np.where(np.isnan(actions), 0, actions)
        self._h_skip.numpy()[:] = skip
        np.copyto(self._h_act.numpy(), actions, casting='unsafe')
        self._d_act.copy_(self._h_act, non_blocking=True)
        self._d_skip.copy_(self._h_skip, non_blocking=True)
        env.step(self._d_act, skip=self._d_skip)
        self._fetch()
        done = self._h_done.numpy().astype(bool)
        stepped = ~skip
        info = self._info
        for i in np.nonzero(stepped & ~done & self._info_dirty)[0]:
            info[i] = {}
        self._info_dirty[stepped] = done[stepped]
        idx_done = np.nonzero(stepped & done)[0]
        if len(idx_done):
            sel = torch.from_numpy(idx_done).to(env.device)
            tobs = env.tobs_out().reshape(env.E, -1, 204)[sel].cpu().numpy()
            rank = self._h_rank.numpy()
            for k, i in enumerate(idx_done):
                info[i] = {'team_ranking': rank[i].astype(np.int64).copy(), 'end_reason': 'EndGame', 'obs-echo': tobs[k]}
        self._suffer_reset[stepped] = done[stepped]
        return self._h_obs.numpy(), self._h_rew.numpy().astype(np.float64), done, info.copy()

    def sleep(self):
        return None
