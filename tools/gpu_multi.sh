#!/bin/bash
# GPU box with N >= 2 GPUs: multi-GPU parity tests, then the slab bench under torchrun
N=${1:-2}
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_multi_$N.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; cat gpurun_out/bench_n$N.json | cut -c1-1500; tail -5 gpurun_out/bench_n$N.err
