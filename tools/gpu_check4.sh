#!/bin/bash
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_gpu.log
timeout 100 build/parallel tests/params/c1.in 2>&1 | tr '\n' ' '; echo
timeout 300 python bench.py --steps 3 --warmup 3 --max-it 20 --no-e2e --no-cpu-baseline 2>&1 | tail -n 1 | cut -c1-400
