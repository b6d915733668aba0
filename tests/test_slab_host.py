"""Host-side logic of the y-slab decomposition, on the CPU: the partition arithmetic and the descriptor
exchange over torch.distributed (gloo, world_size 2).  No compute calls (there is no GPU here)."""
import os
import subprocess
import sys

import pytest

from util import ROOT, nsb


@pytest.mark.parametrize("j_max,chunk,world", [(8192, 120, 1), (8192, 120, 2), (8192, 120, 8), (16384, 120, 4),
                                               (65536, 120, 8), (500, 120, 3), (96, 16, 4), (130, 16, 3)])
def test_partition_covers_the_domain_in_whole_bands(j_max, chunk, world):
    m = nsb()
    parts = [m.slab_partition(j_max, chunk, world, r) for r in range(world)]
    nb = (j_max + chunk - 1) // chunk
    assert all(p[3] == nb for p in parts)
    assert parts[0][0] == 0 and parts[0][2] == 0
    for r in range(world):
        j_lo, nj, band0, _ = parts[r]
        assert j_lo == band0 * chunk and j_lo % 2 == 0              # colour parity survives the cut
        assert nj >= 1
        if r + 1 < world:
            assert parts[r + 1][0] == j_lo + nj and nj % chunk == 0  # only the last slab may end in a short band
    assert parts[-1][0] + parts[-1][1] == j_max
    counts = [(p[1] + chunk - 1) // chunk for p in parts]
    assert max(counts) - min(counts) <= 1 and sum(counts) == nb


def test_partition_rejects_bad_requests():
    m = nsb()
    for args in [(32, 16, 4, 0), (32, 15, 1, 0), (32, 16, 2, 2), (0, 16, 1, 0)]:
        with pytest.raises(m.NsbError):
            m.slab_partition(*args)


WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import importlib
import torch.distributed as dist
m = importlib.import_module("navierstokes-parallel_b200")
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
n = m.lib().nsb_slab_desc_bytes()
gather = m.torch_all_gather()
blobs = gather(bytes([rank + 1]) * n)                 # stand-in descriptor: n bytes of (rank + 1)
assert len(blobs) == world and all(len(b) == n for b in blobs)
assert [b[0] for b in blobs] == [r + 1 for r in range(world)], "descriptors must arrive in rank order"
parts = [m.slab_partition(8192, 120, world, r) for r in range(world)]
mine = parts[rank]
if rank + 1 < world:
    assert parts[rank + 1][0] == mine[0] + mine[1]
# without a GPU the slab itself must refuse to exist (no CPU path)
try:
    m.SlabSolver(m.make_params(i_max=64, j_max=480), rank, world, gather)
    ok = False
except m.NsbError as e:
    ok = "CUDA" in str(e) or "no usable" in str(e)
import torch
if not torch.cuda.is_available():
    assert ok
dist.barrier()
dist.destroy_process_group()
print("worker", rank, "ok")
'''


def test_descriptor_exchange_over_gloo_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29631", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script), ROOT], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert f"worker {r} ok" in o
