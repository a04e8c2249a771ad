"""Box ceiling for the host-facing path (VERDICT r1 "next" #8): N ranks, each copying the 158 MB observation block of one
DcaSuperpoolEnv.step from its GPU into pinned host memory, all at once -- what the PCIe links / the host memory system
give when nothing else runs.  e2e through DcaSuperpoolEnv is bounded by this (the reference surface hands the policy
numpy observations, 157 MB per step per 4096 envs).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/d2h_ceiling.py
prints one JSON line (rank 0): per-rank and aggregate GB/s, max over ranks of the wall time per copy.
"""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    E, G = 4096, 50
    dev = torch.device('cuda', local)
    src = torch.rand((E, G, 192), dtype=torch.float32, device=dev)
    dst = [torch.zeros((E, G, 192), dtype=torch.float32, pin_memory=True) for _ in range(2)]
    nbytes = src.numel() * 4
    for i in range(3):
        dst[i % 2].copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    n = 20
    t0 = time.perf_counter()
    for i in range(n):
        dst[i % 2].copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()   # one copy in flight per rank, like one env step
    dt = (time.perf_counter() - t0) / n
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    try:
        numa = open('/sys/bus/pci/devices/%s/numa_node' % torch.cuda.get_device_properties(local).pci_bus_id.lower()).read().strip()
    except Exception:
        numa = None
    if rank == 0:
        worst = float(t.item())
        print(json.dumps({'what': 'concurrent pinned D2H of one step\'s observations per rank', 'n_gpus': world,
                          'bytes_per_copy': nbytes, 'seconds_per_copy_max_over_ranks': worst,
                          'gbs_per_rank': nbytes / worst / 1e9, 'gbs_aggregate': world * nbytes / worst / 1e9,
                          'e2e_ceiling_agent_steps_per_sec': world * E * 100 / worst,
                          'host_cores': os.cpu_count(), 'gpu0_numa_node': numa}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
