#!/bin/bash
# On the GPU box: kernel-only bench of every build/variants/lib_*.so (plus a quick parity check each).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for so in build/variants/lib_*.so; do
  n=$(basename $so .so)
  NSB200_LIB=$PWD/$so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "per_iteration or committed or scale" > gpurun_out/sweep_$n.test 2>&1
  echo "$n parity: $(tail -1 gpurun_out/sweep_$n.test)"
  NSB200_LIB=$PWD/$so timeout 300 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline "$@" > gpurun_out/sweep_$n.json 2> gpurun_out/sweep_$n.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/sweep_$n.json"))
    print("$n", "value %.0f MLUPS" % d["value"], "kernel %.1f us" % (d["roofline"]["launch_ms"]*1e3), "frac %.3f" % d["roofline"]["frac"], d["clocks"])
except Exception as e:
    print("$n failed", e)
PY
done
