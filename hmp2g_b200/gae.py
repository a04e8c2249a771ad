"""GAE / return over a device rollout buffer: the batched replacement of trajectory.finalize
(/root/reference/ALGORITHM/conc/trajectory.py:60-117, ALGORITHM/common/traj_gae.py:54-133) and of the advantage
normalisation in the PPO sampler (/root/reference/ALGORITHM/conc/ppo.py:53-55)."""
import ctypes as C

import torch

from . import _lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def gae_return(reward, value, dead, traj_len=None, gamma=0.99, tau=0.95, threat_limit=11, want_pushed=False):
    """reward f32 [T,E,G], value f32 [T,E,G], dead uint8/bool [T,E,G] (obs all-NaN, trajectory.py:72),
    traj_len int32 [E] or None.  Returns (ret f32, threat f32[, reward_pushed f32]) -- rows >= traj_len untouched (0)."""
    lib = _lib.load()
    assert reward.is_cuda and reward.dtype == torch.float32 and reward.is_contiguous()
    T, E, G = reward.shape
    assert value.shape == reward.shape and value.dtype == torch.float32 and value.is_contiguous()
    dead = dead.to(torch.uint8).contiguous()
    assert dead.shape == reward.shape
    if traj_len is not None:
        assert traj_len.dtype == torch.int32 and traj_len.shape == (E,) and traj_len.is_cuda
    # full-length trajectories overwrite every element; ragged ones leave rows >= traj_len at zero
    alloc = torch.empty_like if traj_len is None else torch.zeros_like
    ret = alloc(reward)
    threat = alloc(reward)
    pushed = alloc(reward) if want_pushed else None
    stream = C.c_void_p(torch.cuda.current_stream(reward.device).cuda_stream)
    with torch.cuda.device(reward.device):
        _lib.check(lib.dca_gae(_ptr(reward), _ptr(value), _ptr(dead), _ptr(traj_len), _ptr(ret), _ptr(threat),
                               _ptr(pushed), T, E, G, float(gamma), float(tau), int(threat_limit), stream))
    return (ret, threat, pushed) if want_pushed else (ret, threat)


def normalized_advantage(ret, value, valid=None, group=None):
    """adv = ret - value; adv = (adv - mean) / (std + 1e-5) over the pooled batch (ppo.py:53-55).
    With `group` (a torch.distributed process group) the moments are all-reduced over ranks (NCCL): the only
    collective on the rollout path besides the gradient all-reduce."""
    adv = ret - value
    if valid is None:
        valid = torch.ones_like(adv, dtype=torch.bool)
    m = valid.to(adv.dtype)
    stats = torch.stack([(adv * m).sum(), (adv * adv * m).sum(), m.sum()]).to(torch.float64)
    if group is not None or (torch.distributed.is_available() and torch.distributed.is_initialized()):
        torch.distributed.all_reduce(stats, group=group)
    n = stats[2]
    mean = stats[0] / n
    var = (stats[1] / n - mean * mean).clamp_min(0) * (n / (n - 1).clamp_min(1))  # torch.std default: unbiased
    return ((adv - mean.to(adv.dtype)) / (var.sqrt().to(adv.dtype) + 1e-5)) * m
