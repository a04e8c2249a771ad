"""Diagnostics (needs a -DDCA_DIAG_COUNT build in tools/variants/diagcount.so): kNN close-call statistics per step."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmp2g_b200.config import ScenarioConfig  # noqa: E402
from hmp2g_b200.vec_env import DcaVecEnv  # noqa: E402

E = 4096
scfg = type('S', (ScenarioConfig,), dict(bench.WORKLOAD))
env = DcaVecEnv(E, scfg=scfg, seed=1)
env.reset()
g = torch.Generator(device=env.device)
g.manual_seed(0)
out = (C.c_ulonglong * 8)()
env.lib.dca_debug_counters(out, 1)
for t in range(180):
    act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
    env.step(act, env.plan_centroid())
    if t % 15 == 14:
        env.lib.dca_debug_counters(out, 1)
        v = np.array(list(out), dtype=np.float64) / 15 / E
        print('steps %3d-%3d: tasks/env %.1f  close calls/env %.2f  candidates per close call %.1f  same-bucket %.2f  alive %.2f'
              % (t - 14, t, v[0], v[1], v[2] / max(v[1], 1e-9), v[3], float(env.alive.float().mean())))
