#!/bin/bash
# Run on the GPU box (gpurun): launch list + full ncu capture of the dominant kernels.  $1 = tag (e.g. r1d)
TAG=${1:-prof}
CMD="python bench.py --steps 2 --warmup 1 --max-it 20 --no-e2e --no-cpu-baseline --no-parity"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sor_persist -s 2 -c 2 -o gpurun_out/${TAG}_sor_persist $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'fg_kernel|rhs_kernel|update_uv|fgrhs' -s 2 -c 4 -o gpurun_out/${TAG}_momentum $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
for f in plain ncu1 ncu2 ncu3; do tail -n 3 gpurun_out/${TAG}_$f.log; done
