"""Device-resident rollout buffer feeding the GAE kernel: the zero-copy replacement of the trajectory pool of
/root/reference/ALGORITHM/conc/trajectory.py (BatchTrajManager / trajectory.finalize :67-117) for envs that run on the
GPU.  Layout [T, E, G] is the one dca_gae scans (time-major, (env, guard) columns coalesced).

One trajectory per env per episode, as in the reference: when an env finishes an episode its column is finalized
(reward push-forward, threat, GAE, return) over the steps it collected and handed out; the column restarts at t = 0.
`dead[t, e, g]` = "the observation guard g acted on at step t was all-NaN" (trajectory.py:72), i.e. the guard was not
alive when the step began.
"""
import torch

from .gae import gae_return, normalized_advantage


class DeviceRollout(object):
    def __init__(self, env, horizon=None, gamma=0.99, tau=0.95, threat_limit=11):
        self.env = env
        E, G = env.E, env.G
        self.T = int(horizon or env.cfg.max_episode_step)
        dev = env.device
        self.reward = torch.zeros((self.T, E, G), dtype=torch.float32, device=dev)
        self.value = torch.zeros((self.T, E, G), dtype=torch.float32, device=dev)
        self.dead = torch.zeros((self.T, E, G), dtype=torch.uint8, device=dev)
        self.length = torch.zeros((E,), dtype=torch.int32, device=dev)   # steps collected in the running episode
        self.gamma, self.tau, self.threat_limit = gamma, tau, threat_limit
        self._rows = torch.arange(E, device=dev)

    def alive_guards(self):
        """uint8 [E,G]: who is alive in the observation the policy is about to act on."""
        return self.env.alive[:, :self.env.G]

    def add(self, alive_before, value, reward, done):
        """Record one env step.  alive_before: alive_guards() taken BEFORE env.step; value f32 [E,G] = critic output on
        that observation; reward f32 [E,G], done u8 [E] = outputs of env.step.  Returns the mask of finished envs."""
        t = self.length.long().clamp_max(self.T - 1)
        self.reward[t, self._rows] = reward
        self.value[t, self._rows] = value
        self.dead[t, self._rows] = (alive_before == 0).to(torch.uint8)
        self.length += 1
        return done != 0

    def finalize(self, finished=None, group=None):
        """trajectory.finalize for the envs in `finished` (bool [E]; None = all, e.g. a truncated rollout): returns
        (ret, threat, adv, valid) over the whole buffer, `valid[t, e]` marking the steps that belong to the finalized
        trajectories; their columns restart."""
        E = self.env.E
        if finished is None:
            finished = torch.ones((E,), dtype=torch.bool, device=self.reward.device)
        lens = torch.where(finished, self.length, torch.zeros_like(self.length)).clamp_max(self.T)
        ret, threat = gae_return(self.reward, self.value, self.dead, traj_len=lens, gamma=self.gamma, tau=self.tau,
                                 threat_limit=self.threat_limit)
        valid = (torch.arange(self.T, device=lens.device)[:, None] < lens[None, :])[:, :, None].expand_as(ret)
        adv = normalized_advantage(ret, self.value, valid=valid, group=group)
        self.length = torch.where(finished, torch.zeros_like(self.length), self.length)
        return ret, threat, adv, valid
