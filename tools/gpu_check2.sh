#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "resident" > gpurun_out/pytest_res.log 2>&1; echo "resident tests rc=$?"; tail -n 6 gpurun_out/pytest_res.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3
timeout 200 bash tools/ref_cuda_compare.sh 1 2>&1 | tail -n 4
timeout 400 python -m pytest tests/test_gpu_ghia.py -m gpu -q -s > gpurun_out/pytest_ghia.log 2>&1; echo "ghia rc=$?"; grep -E "Re=|passed|failed|Error" gpurun_out/pytest_ghia.log | head -n 12
