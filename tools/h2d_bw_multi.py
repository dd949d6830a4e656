"""Concurrent pinned H2D bandwidth of all ranks, without and with NUMA binding:
python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/h2d_bw_multi.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from opticalflow_b200.driver import bind_host_to_gpu
rank, lr, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl")
n = 88080384
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def run(tag):
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); h.fill_(1)
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30): d.copy_(h, non_blocking=True)
    e1.record(); e1.synchronize()
    t = e0.elapsed_time(e1) / 30 * 1e-3
    bw = torch.tensor([n / t / 1e9], device="cuda")
    if world > 1:
        allbw = [torch.zeros_like(bw) for _ in range(world)]; dist.all_gather(allbw, bw)
    else: allbw = [bw]
    if rank == 0:
        v = [round(float(b), 1) for b in allbw]
        print(f"{tag}: per-rank GB/s {v} sum {sum(v):.0f}", flush=True)
print(f"rank {rank}: allowed cpus {len(os.sched_getaffinity(0))} of {os.cpu_count()}", flush=True)
run("unbound")
cpus = bind_host_to_gpu(lr)
print(f"rank {rank}: bound to {None if cpus is None else (len(cpus), cpus[0], cpus[-1])}", flush=True)
run("bound  ")
