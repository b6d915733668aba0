#!/bin/bash
# GPU box: guarded test + probe run (every command under its own timeout, output written as it goes)
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 120 python tools/wave_probe.py 120 13,26,69,103 2>&1 | tee gpurun_out/wave_probe2.txt | cut -c1-150
echo "probe rc=$?"
