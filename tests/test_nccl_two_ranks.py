"""SURVEY.md section 8(e) on hardware: the three collectives of the multi-GPU path over NCCL, world_size 2 on one box --
rollout statistics, advantage moments (dca_adv_moments + all-reduce of 3 doubles), DynamicNorm moments (dca_norm_moments +
all-reduce of 49 doubles) -- and that the env shards by global env id (rank r's envs equal the same global ids of a
single-rank run).  Needs two GPUs: skipped on a one-GPU box (the gloo twin in test_host_cpu.py covers the host logic)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r'''
import os, sys, torch
sys.path.insert(0, %r)
import torch.distributed as dist
from hmp2g_b200.sharding import init_distributed, shard_envs, RolloutStats
from hmp2g_b200.gae import normalized_advantage
from hmp2g_b200.policy import DynamicNormDevice
from hmp2g_b200.config import ScenarioConfig
from hmp2g_b200.vec_env import DcaVecEnv
rank, world, local = init_distributed('nccl')
dev = torch.device('cuda', local)
torch.cuda.set_device(dev)
E, G = 64, 50
gid0, n = shard_envs(E, rank, world)
g = torch.Generator().manual_seed(0)
ret = torch.randn(30, E, G, generator=g); val = torch.randn(30, E, G, generator=g)
# advantage normalisation over the pooled batch of both ranks == the single-process formula (ppo.py:53-55)
adv = normalized_advantage(ret[:, gid0:gid0 + n].contiguous().to(dev), val[:, gid0:gid0 + n].contiguous().to(dev))
full = (ret - val); want = (full - full.mean()) / (full.std(unbiased=False) + 1e-5)
assert torch.allclose(adv.cpu(), want[:, gid0:gid0 + n], rtol=1e-5, atol=1e-5)
# DynamicNorm: both ranks end with the moments of the whole batch
x = torch.randn(E, G, 24, 16, generator=g); x[torch.rand(E, G, 24, generator=g) < 0.3] = float('nan')
bn = DynamicNormDevice(16).to(dev).train()
bn.update(x[gid0:gid0 + n].contiguous().to(dev))
rows = x.reshape(-1, 16); rows = rows[~torch.isnan(rows).any(-1)].double()
assert int(bn.n_sample.item()) == rows.shape[0]
assert torch.allclose(bn.mean.cpu().double(), rows.mean(0), rtol=1e-5, atol=1e-6)
assert torch.allclose(bn.var.cpu().double(), rows.var(0, unbiased=False).clamp(0.01, 1000), rtol=1e-5, atol=1e-6)
# rollout statistics
reward = torch.randn(E, G, generator=g); done = (torch.rand(E, generator=g) < 0.5).to(torch.uint8)
win = (torch.rand(E, generator=g) < 0.5).to(torch.uint8)
st = RolloutStats(dev); sl = slice(gid0, gid0 + n)
st.update(reward[sl].to(dev), done[sl].to(dev), win[sl].to(dev))
tot = st.reduce()
ref = RolloutStats('cpu'); ref.update(reward, done, win)
for k, v in zip(ref.FIELDS, ref.buf.tolist()):
    assert abs(v - tot[k]) <= 1e-4 * max(1.0, abs(v)), (k, v, tot[k])  # reward sums are float32 reductions on the device
# env sharding: this rank's slice of a 2-rank run == the same global env ids in a run that holds all of them
scfg = type('S', (ScenarioConfig,), dict(num_guards=50, num_attackers=50, MaxEpisodeStep=30))
mine = DcaVecEnv(n, scfg=scfg, device=dev, seed=9, env_gid0=gid0)
allenvs = DcaVecEnv(E, scfg=scfg, device=dev, seed=9, env_gid0=0)
mine.reset(); allenvs.reset()
ga = torch.Generator().manual_seed(3)
for t in range(40):
    a = torch.randint(0, 7, (E, G), generator=ga, dtype=torch.int32)
    mine.step(a[gid0:gid0 + n].contiguous().to(dev)); allenvs.step(a.to(dev))
torch.cuda.synchronize()
assert torch.equal(mine.obs, allenvs.obs[gid0:gid0 + n]) and torch.equal(mine.reward, allenvs.reward[gid0:gid0 + n])
assert torch.equal(mine.done, allenvs.done[gid0:gid0 + n]) and torch.equal(mine.alive, allenvs.alive[gid0:gid0 + n])
dist.barrier(); dist.destroy_process_group()
sys.stdout.write('rank' + str(rank) + 'ok\n'); sys.stdout.flush()
'''


def test_two_rank_nccl_collectives_and_env_sharding(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (gpurun --gpus 2); one-GPU boxes run the gloo twin in test_host_cpu.py')
    script = tmp_path / 'worker.py'
    script.write_text(_WORKER % REPO)
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', OMP_NUM_THREADS='1')
    port = 29700 + os.getpid() % 200
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                          '--master-addr', '127.0.0.1', '--master-port', str(port), str(script)],
                         env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert out.stdout.count('ok') == 2, out.stdout
