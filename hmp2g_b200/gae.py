"""GAE / return over a device rollout buffer: the batched replacement of trajectory.finalize
(/root/reference/ALGORITHM/conc/trajectory.py:60-117, ALGORITHM/common/traj_gae.py:54-133) and of the advantage
normalisation in the PPO sampler (/root/reference/ALGORITHM/conc/ppo.py:53-55)."""
import ctypes as C

import torch

from . import _lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def gae_return(reward, value, dead, traj_len=None, gamma=0.99, tau=0.95, threat_limit=11, want_pushed=False,
               gamma_fwd=None):
    """reward f32 [T,E,G], value f32 [T,E,G], dead uint8/bool [T,E,G] (obs all-NaN, trajectory.py:72),
    traj_len int32 [E] or None.  gamma_fwd: AlgorithmConfig.gamma_in_reward_forwarding_value when
    gamma_in_reward_forwarding is on (traj_gae.py:55-60), None = off.
    Returns (ret f32, threat f32[, reward_pushed f32]) -- rows >= traj_len untouched (0)."""
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
                               _ptr(pushed), T, E, G, float(gamma), float(tau),
                               1.0 if gamma_fwd is None else float(gamma_fwd), int(threat_limit), stream))
    return (ret, threat, pushed) if want_pushed else (ret, threat)


def gae_return_levels(reward, value_levels, dead, traj_len=None, gamma=0.99, tau=0.95, threat_limit=8, gamma_fwd=None):
    """The policy-resonance variant of finalize (ALGORITHM/common/traj_gae.py:86-90, gae_finalize_return_pr :135-169):
    the critic emits `distribution_precision` value levels per agent (value_levels f32 [T,E,G,P]); the reward is pushed
    forward and the threat counted once per agent, then the same GAE scan runs on every level, fed by the agent's
    pushed reward.  Two dca_gae launches: [T,E,G] for push-forward / threat, [T,E,G*P] (no dead frames) for the returns.
    Returns (return_levels f32 [T,E,G,P], threat f32 [T,E,G], reward_pushed f32 [T,E,G])."""
    T, E, G, P = value_levels.shape
    assert reward.shape == (T, E, G)
    v0 = value_levels[..., 0].contiguous()
    _, threat, pushed = gae_return(reward, v0, dead, traj_len=traj_len, gamma=gamma, tau=tau, threat_limit=threat_limit,
                                   want_pushed=True, gamma_fwd=gamma_fwd)
    rep = pushed.unsqueeze(-1).expand(T, E, G, P).reshape(T, E, G * P).contiguous()
    nodead = torch.zeros((T, E, G * P), dtype=torch.uint8, device=reward.device)
    ret, _ = gae_return(rep, value_levels.reshape(T, E, G * P).contiguous(), nodead, traj_len=traj_len, gamma=gamma, tau=tau,
                        threat_limit=threat_limit)
    return ret.reshape(T, E, G, P), threat, pushed


def allreduce_moments(stats, group=None):
    """stats f64[3] = (sum, sum of squares, count) of this rank's advantages -> the same over all ranks (in place)."""
    if group is not None or (torch.distributed.is_available() and torch.distributed.is_initialized()):
        torch.distributed.all_reduce(stats, group=group)
    return stats


def moments_mean_std(stats):
    """(mean, population std) from (sum, sum of squares, count) -- what k_adv_normalize derives on the device."""
    n = stats[2].clamp_min(1)
    mean = stats[0] / n
    return mean, (stats[1] / n - mean * mean).clamp_min(0).sqrt()


def normalized_advantage(ret, value, valid=None, group=None):
    """adv = ret - value; adv = (adv - mean) / (std + 1e-5) over the pooled batch (ppo.py:53-55; numpy std, ddof 0),
    by dca_adv_moments / dca_adv_normalize.  valid: bool / uint8 mask, either elementwise (ret.shape) or one flag per
    leading index (ret.shape[:-1], e.g. [T,E] for ret [T,E,G]); invalid elements get 0.  If torch.distributed is
    initialised (or `group` is given) the three moments are all-reduced over the ranks (NCCL): the only collective on
    the rollout path besides the gradient all-reduce."""
    lib = _lib.load()
    assert ret.is_cuda and ret.dtype == torch.float32 and ret.is_contiguous()
    assert value.shape == ret.shape and value.dtype == torch.float32 and value.is_contiguous()
    n = ret.numel()
    vgroup = 1
    if valid is not None:
        if tuple(valid.shape) == tuple(ret.shape[:-1]):
            vgroup = ret.shape[-1]
        else:
            assert valid.shape == ret.shape, 'valid must match ret or ret.shape[:-1]'
        valid = valid.to(torch.uint8).contiguous()
    dev = ret.device
    scratch = torch.empty((592 * 3,), dtype=torch.float64, device=dev)
    stats = torch.empty((3,), dtype=torch.float64, device=dev)
    adv = torch.empty_like(ret)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    with torch.cuda.device(dev):
        _lib.check(lib.dca_adv_moments(_ptr(ret), _ptr(value), _ptr(valid), vgroup, n, _ptr(scratch), _ptr(stats),
                                       stream))
        allreduce_moments(stats, group)
        _lib.check(lib.dca_adv_normalize(_ptr(ret), _ptr(value), _ptr(valid), vgroup, n, _ptr(stats), _ptr(adv),
                                         stream))
    return adv
