"""Aggregate pinned host->device copy rate with one rank per GPU (torchrun).

Explains the multi-GPU end-to-end line of bench.py: that leg uploads every step's token rows
from pinned host memory, so its ceiling at N ranks is whatever the host platform delivers to
N GPUs at once, which this measures without any of the repo's kernels in the way.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29561 tools/h2d_scaling.py [--bind] [--mb 210]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--mb', type=int, default=210, help='bytes per copy (bench step uploads ~210 MB)')
    ap.add_argument('--iters', type=int, default=20)
    ap.add_argument('--bind', action='store_true', help='bind to the GPU-local CPUs first')
    a = ap.parse_args()

    rank = int(os.environ.get('RANK', 0))
    local = int(os.environ.get('LOCAL_RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    bound = 0
    out = os.fdopen(os.dup(1), 'w')      # bench.py points fd 1 at stderr when imported
    if a.bind:
        from bench import bind_to_gpu_numa_node
        bound = bind_to_gpu_numa_node(local)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    n = a.mb << 20
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    dev = torch.empty(n, dtype=torch.uint8, device='cuda')
    for _ in range(3):
        dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        dev.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    gbs = torch.tensor([n * a.iters / ms / 1e6], device='cuda', dtype=torch.float64)
    if world > 1:
        all_gbs = [torch.zeros_like(gbs) for _ in range(world)]
        dist.all_gather(all_gbs, gbs)
        rates = [float(t) for t in all_gbs]
    else:
        rates = [float(gbs)]
    if rank == 0:
        print(json.dumps({'n_gpus': world, 'bind': bool(a.bind), 'cpus_bound_rank0': bound,
                          'mb_per_copy': a.mb, 'gbs_per_rank': [round(r, 2) for r in rates],
                          'gbs_total': round(sum(rates), 2)}), file=out, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
