"""Device-resident rollout buffer feeding the GAE kernel: the zero-copy replacement of the trajectory pool of
/root/reference/ALGORITHM/conc/trajectory.py (BatchTrajManager / trajectory.finalize :67-117) for envs that run on the
GPU.  Layout [T, E, G] is the one dca_gae scans (time-major, (env, guard) columns coalesced).

One trajectory per env per episode, as in the reference: when an env finishes an episode its column is finalized
(reward push-forward, threat, GAE, return) over the steps it collected and handed out; the column restarts at t = 0.
`dead[t, e, g]` = "the observation guard g acted on at step t was all-NaN" (trajectory.py:72), i.e. the guard was not
alive when the step began.
"""
import ctypes as C

import torch

from . import _lib
from .gae import gae_return, normalized_advantage


class HistoryPool(object):
    """The conc_4hist history pool on the device: ShellEnvWrapper's '_Histpool_Obs_' state and solve_duplicate
    (/root/reference/ALGORITHM/conc_4hist/shell_env.py:25-99, roll_hisory cython_func.pyx:32-53).  roll() takes the
    env's observation tensor [E,G,12,16] (zero-copy from DcaVecEnv.obs) and returns obs_feed [E,G,24,16]: the new rows
    followed by the remembered ones, empty rows NaN, everything NaN for a dead agent or a paused env."""

    def __init__(self, n_env, n_agent, device=None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.DcaError('hmp2g_b200 needs a CUDA device; there is no CPU fallback')
        self.device = torch.device(device if device is not None else 'cuda')
        assert self.device.type == 'cuda'
        if self.device.index is None:  # 'cuda' means the current device, exactly as DcaVecEnv resolves it
            self.device = torch.device('cuda', torch.cuda.current_device())
        self.E, self.G = int(n_env), int(n_agent)
        self.his_pool = torch.zeros((self.E, self.G, 12, 16), dtype=torch.float32, device=self.device)
        self.obs_feed = torch.zeros((self.E, self.G, 24, 16), dtype=torch.float32, device=self.device)

    def roll(self, obs, pause=None, suffered_reset=None):
        if obs.dtype != torch.float32 or obs.numel() != self.his_pool.numel() or not obs.is_cuda \
                or not obs.is_contiguous():
            raise _lib.DcaError('roll: obs must be a contiguous cuda float32 [E,G,12,16]')
        for m in (pause, suffered_reset):
            if m is not None and (m.dtype != torch.uint8 or tuple(m.shape) != (self.E,) or not m.is_cuda):
                raise _lib.DcaError('roll: masks must be cuda uint8 [E]')
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        _lib.check(self.lib.dca_histpool_roll(p(obs), p(self.his_pool), p(pause), p(suffered_reset), p(self.obs_feed),
                                              self.E, self.G,
                                              C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return self.obs_feed


class DeviceRollout(object):
    def __init__(self, env, horizon=None, gamma=0.99, tau=0.95, threat_limit=11, gamma_fwd=None):
        self.env = env
        E, G = env.E, env.G
        self.T = int(horizon or env.cfg.max_episode_step)
        dev = env.device
        self.reward = torch.zeros((self.T, E, G), dtype=torch.float32, device=dev)
        self.value = torch.zeros((self.T, E, G), dtype=torch.float32, device=dev)
        self.dead = torch.zeros((self.T, E, G), dtype=torch.uint8, device=dev)
        self.length = torch.zeros((E,), dtype=torch.int32, device=dev)   # steps collected in the running episode
        self.gamma, self.tau, self.threat_limit = gamma, tau, threat_limit
        self.gamma_fwd = gamma_fwd  # gamma_in_reward_forwarding_value, None = the flag is off (traj_gae.py:54-60)
        self._rows = torch.arange(E, device=dev)

    def alive_guards(self):
        """uint8 [E,G]: who is alive in the observation the policy is about to act on."""
        return self.env.alive[:, :self.env.G]

    def add(self, alive_before, value, reward, done, active=None):
        """Record one env step.  alive_before: alive_guards() taken BEFORE env.step; value f32 [E,G] = critic output on
        that observation; reward f32 [E,G], done u8 [E] = outputs of env.step; active (bool / u8 [E], None = all):
        the envs that really stepped (~ENV-PAUSE) -- paused envs record nothing, like the reference, which feeds
        only obs[~P] to the algorithm (ALGORITHM/conc/shell_env.py:44-50).  Returns the mask of finished envs."""
        t = self.length.long().clamp_max(self.T - 1)
        if active is None:
            self.reward[t, self._rows] = reward
            self.value[t, self._rows] = value
            self.dead[t, self._rows] = (alive_before == 0).to(torch.uint8)
            self.length += 1
            return done != 0
        act = active != 0
        rows, tt = self._rows[act], t[act]
        self.reward[tt, rows] = reward[act]
        self.value[tt, rows] = value[act]
        self.dead[tt, rows] = (alive_before[act] == 0).to(torch.uint8)
        self.length += act.to(self.length.dtype)
        return (done != 0) & act

    def finalize(self, finished=None, group=None):
        """trajectory.finalize for the envs in `finished` (bool [E]; None = all, e.g. a truncated rollout): returns
        (ret, threat, adv, valid) over the whole buffer, `valid[t, e]` marking the steps that belong to the finalized
        trajectories; their columns restart."""
        E = self.env.E
        if finished is None:
            finished = torch.ones((E,), dtype=torch.bool, device=self.reward.device)
        lens = torch.where(finished, self.length, torch.zeros_like(self.length)).clamp_max(self.T)
        ret, threat = gae_return(self.reward, self.value, self.dead, traj_len=lens, gamma=self.gamma, tau=self.tau,
                                 threat_limit=self.threat_limit, gamma_fwd=self.gamma_fwd)
        valid_te = torch.arange(self.T, device=lens.device)[:, None] < lens[None, :]
        adv = normalized_advantage(ret, self.value, valid=valid_te, group=group)
        valid = valid_te[:, :, None].expand_as(ret)
        self.length.copy_(torch.where(finished, torch.zeros_like(self.length), self.length))  # in place (CUDA graphs)
        return ret, threat, adv, valid
