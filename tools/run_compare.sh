#!/bin/bash
# The reference's benchmark procedure (run.sh / run_parallel_block_tests.sh / colab-runner.ipynb) against
# this build, on the GPU box: for every input run the reference serial binary (optional, slow), the
# reference CUDA binary (oracle/_ref/parallel, block 16) and build/parallel; compare the stdout with the
# notebook's tolerance rule; write results/speedup.csv in the layout of the reference's speedup.csv
# (test, serial SOR seconds, reference CUDA SOR seconds, nsb200 SOR seconds, speed-ups).
#   tools/run_compare.sh [--serial] 1 2 3
cd "$(dirname "$0")/.."
SERIAL=0; [ "$1" = "--serial" ] && { SERIAL=1; shift; }
OUT=${NSB_RESULTS:-gpurun_out/results}; mkdir -p $OUT
echo "test,grid,serial_sor_s,reference_cuda_sor_s,nsb200_sor_s,nsb200_vs_reference_cuda,nsb200_vs_serial,stdout_check" > $OUT/speedup.csv
for k in ${@:-1 2}; do
  f=tests/params/c$k.in; n=$(sed -n 3p $f | awk '{print $1}')
  ts=""; tr=""; chk="n/a"
  if [ $SERIAL = 1 ] && [ -x oracle/_ref/serial ]; then timeout 3000 oracle/_ref/serial $f > $OUT/serial_$k.out 2> $OUT/serial_$k.err && ts=$(grep -oE '[0-9]+\.[0-9]+' $OUT/serial_$k.err | tail -n 1); fi
  if [ -x oracle/_ref/parallel ]; then timeout 1500 oracle/_ref/parallel $f 16 > $OUT/refcuda_$k.out 2> $OUT/refcuda_$k.err && tr=$(grep -oE '[0-9]+\.[0-9]+' $OUT/refcuda_$k.err | tail -n 1); fi
  timeout 1500 build/parallel $f 16 > $OUT/nsb_$k.out 2> $OUT/nsb_$k.err; to=$(grep -oE '[0-9]+\.[0-9]+' $OUT/nsb_$k.err | tail -n 1)
  [ -n "$tr" ] && chk=$(python tools/compare_outputs.py $OUT/refcuda_$k.out $OUT/nsb_$k.out | cut -d: -f1)
  [ -n "$ts" ] && chk="$chk/$(python tools/compare_outputs.py $OUT/serial_$k.out $OUT/nsb_$k.out | cut -d: -f1)"
  python - "$k" "$n" "$ts" "$tr" "$to" "$chk" >> $OUT/speedup.csv <<'PY'
import sys
k, n, ts, tr, to, chk = sys.argv[1:7]
f = lambda a, b: f"{float(a) / float(b):.2f}" if a and b and float(b) > 0 else ""
print(",".join([k, f"{n}x{n}", ts, tr, to, f(tr, to), f(ts, to), chk]))
PY
  tail -n 1 $OUT/speedup.csv
done
