#!/usr/bin/env python
"""Host <-> device copy bandwidth with N ranks copying AT THE SAME TIME (the ceiling of bench.py's e2e leg at N > 1):
   torchrun --nproc-per-node N tools/h2d_probe.py [--bind 0|1]
Every rank copies what one e2e step copies (3 slabs of 8194 x 8194 doubles, 1.6 GB each way) from / to pinned memory,
H2D alone, D2H alone, then both directions at once on two streams.  Prints per-rank GB/s (min / mean over ranks)."""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--bind", type=int, default=1)
ap.add_argument("--mb", type=int, default=1611)
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
binding = bench.bind_to_gpu_numa(local) if a.bind else {"numa_node": None, "note": "binding off"}
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = a.mb * 1000 * 1000 // 8
h_in = torch.zeros(n, dtype=torch.float64).pin_memory()
h_out = torch.zeros(n, dtype=torch.float64).pin_memory()
d_a = torch.empty(n, dtype=torch.float64, device="cuda")
d_b = torch.ones(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, reps=3):
    fn(); barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    el = (time.perf_counter() - t0) / reps
    barrier()
    return el


def h2d():
    with torch.cuda.stream(s1):
        d_a.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_b, non_blocking=True)


def both():
    h2d(); d2h()


gb = n * 8 / 1e9
res = [gb / timed(h2d), gb / timed(d2h), 2 * gb / timed(both)]
t = torch.tensor(res, device="cuda", dtype=torch.float64)
if world > 1:
    lo, su = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(su, op=dist.ReduceOp.SUM)
else:
    lo, su = t, t
if rank == 0:
    names = ("H2D alone", "D2H alone", "H2D + D2H at once (sum of both directions)")
    print(f"{world} rank(s) copying {gb:.2f} GB each at the same time; rank 0 binding: {binding}; host cores {os.cpu_count()}")
    for k, nm in enumerate(names):
        print(f"  {nm:45s} per rank: min {lo[k].item():6.1f} GB/s, mean {su[k].item() / world:6.1f} GB/s; all ranks together {su[k].item():7.1f} GB/s")
if world > 1:
    dist.destroy_process_group()
