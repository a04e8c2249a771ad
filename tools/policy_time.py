"""ConcNet4Hist.act at 4096 x 50 agents: fused kernels vs the torch path (CUDA events), plus a kernel breakdown."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hmp2g_b200.policy import ConcNet4Hist
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden', 'policy_net_4hist.npz'))
net = ConcNet4Hist(16, 7, 3)
net.load_state_dict({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith('sd/')})
net = net.cuda().train()
x = torch.from_numpy(z['obs_feed']).cuda().repeat(205, 1, 1, 1)[:4096].contiguous()
g = torch.Generator(device='cuda'); g.manual_seed(0)
for fused in (True, False):
    ConcNet4Hist.fused = fused
    for _ in range(3): net.act(x, generator=g)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(10): net.act(x, generator=g)
    b.record(); torch.cuda.synchronize()
    print('fused=%d: act %.3f ms per call (4096 x 50 agents)' % (fused, a.elapsed_time(b) / 10))
ConcNet4Hist.fused = True
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    net.act(x, generator=g); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=14, max_name_column_width=60))
