import sys, os, torch
sys.path.insert(0, '/root/repo')
import bench
from hmp2g_b200.config import ScenarioConfig
from hmp2g_b200.vec_env import DcaVecEnv
scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
env = DcaVecEnv(4096, scfg=scfg, seed=1)
env.reset()
g = torch.Generator(device=env.device); g.manual_seed(0)
for t in range(60):
    act = torch.randint(0, 7, (4096, env.G), generator=g, device=env.device, dtype=torch.int32)
    env.step(act, env.plan())
torch.cuda.synchronize()
print('ok')
