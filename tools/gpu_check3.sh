#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "resident or committed or convergence" > gpurun_out/pytest_res.log 2>&1; echo "resident tests rc=$?"; tail -n 4 gpurun_out/pytest_res.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
for c in 2 4 8 16; do
  echo "cluster >= $c"; NSB_RES_CLUSTER=$c timeout 100 build/parallel tests/params/c1.in 2>&1 | tr '\n' ' '; echo
done
