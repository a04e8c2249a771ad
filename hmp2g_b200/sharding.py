"""Multi-GPU layout: envs shard by env index, one process per GPU, no collective in step / obs / reset / GAE.

The reference has a single parallelism axis -- `num_threads` envs over worker processes
(/root/reference/config.py:74-76, UTIL/shm_pool.pyx:259-283) -- and no collective backend.  Here rank r owns the
contiguous global env ids [gid0, gid0 + n_local); the keyed RNG uses the GLOBAL id, so a run is bit-identical for any
world size.  torch.distributed (NCCL over NVLink on the box, gloo in the CPU tests) is used only for rollout
statistics and advantage moments (and by the learner for its gradient all-reduce).
"""
import os

import torch
import torch.distributed as dist


def shard_envs(total_envs, rank, world_size):
    """Contiguous block partition; the first (total % world) ranks get one extra env."""
    assert 0 <= rank < world_size and total_envs >= world_size
    base, extra = divmod(total_envs, world_size)
    n_local = base + (1 if rank < extra else 0)
    gid0 = rank * base + min(rank, extra)
    return gid0, n_local


def init_distributed(backend=None):
    """Rendezvous from torchrun's env (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*).  Returns (rank, world, local_rank)."""
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29500')
        if backend is None:
            backend = 'nccl' if torch.cuda.is_available() else 'gloo'
        kw = {}
        if backend == 'nccl':
            torch.cuda.set_device(local_rank)
            kw['device_id'] = torch.device('cuda', local_rank)  # binds the communicator to this rank's GPU
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world, local_rank


class RolloutStats(object):
    """Device-side episode counters (episodes, wins, reward sum, agent-steps) reduced with ONE small all-reduce.
    Replaces the O(E) Python loop of task_runner.update_runner (/root/reference/task_runner.py:105-133)."""
    FIELDS = ('episodes', 'wins', 'reward_sum', 'env_steps')

    def __init__(self, device):
        self.buf = torch.zeros(len(self.FIELDS), dtype=torch.float64, device=device)

    def update(self, reward, done, win, skip=None):
        active = torch.ones_like(done, dtype=torch.bool) if skip is None else (skip == 0)
        d = (done != 0) & active
        self.buf[0] += d.sum()
        self.buf[1] += ((win != 0) & d).sum()
        self.buf[2] += (reward.sum(dim=1) * active).sum()
        self.buf[3] += active.sum()

    def reduce(self, group=None):
        out = self.buf.clone()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(out, group=group)
        return dict(zip(self.FIELDS, out.tolist()))
