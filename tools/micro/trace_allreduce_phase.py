import os, sys, time, json
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, ROOT+'/gptq-v2_b200')
import torch, torch.distributed as dist
import gptq, sharded
from gptq_py import HessianHook
rank, local = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
world = dist.get_world_size()
dev = torch.device('cuda', local)
n_local = 128 // world
X = {D: torch.randn(n_local, 2048, D, device=dev).half() for D in (4096, 11008)}
def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e
for it in range(4):
    rows = []
    for D in (4096, 4096, 4096, 11008):
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter(); e0 = ev()
        hook = HessianHook()
        for i in range(0, n_local, 16):
            hook.add_batch(X[D][i:i + 16])
        t1 = time.perf_counter(); e1 = ev()
        packed = gptq.pack_upper(hook.hessian)
        t2 = time.perf_counter(); e2 = ev()
        dist.all_reduce(packed)
        t3 = time.perf_counter(); e3 = ev()
        gptq.unpack_upper(packed, hook.hessian)
        t4 = time.perf_counter(); e4 = ev()
        torch.cuda.synchronize()
        rows.append({'D': D, 'host_ms': [round((b - a) * 1e3, 3) for a, b in ((t0, t1), (t1, t2), (t2, t3), (t3, t4))],
                     'gpu_ms': [round(a.elapsed_time(b), 3) for a, b in ((e0, e1), (e1, e2), (e2, e3), (e3, e4))]})
    if rank == 0 and it >= 2:
        print(json.dumps(rows))
dist.destroy_process_group()
