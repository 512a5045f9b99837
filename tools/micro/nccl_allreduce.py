#!/usr/bin/env python
"""NCCL all-reduce timing of the Hessian payloads (fp32): torchrun --nproc-per-node N tools/micro/nccl_allreduce.py"""
import json
import os
import torch
import torch.distributed as dist

rank, local = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
world = dist.get_world_size()
for name, n in (('packed upper D=4096', 528 * 128 * 128), ('packed upper D=11008', 3741 * 128 * 128), ('full D=11008', 11008 * 11008)):
    t = torch.ones(n, device='cuda')
    for _ in range(3):
        dist.all_reduce(t)
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        dist.all_reduce(t)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    if rank == 0:
        print(json.dumps({'k': 'nccl_allreduce', 'what': name, 'world': world, 'MB': n * 4 / 1e6, 'ms': ms,
                          'algbw_GBps': n * 4 / ms / 1e6, 'busbw_GBps': n * 4 / ms / 1e6 * 2 * (world - 1) / world}))
dist.destroy_process_group()
