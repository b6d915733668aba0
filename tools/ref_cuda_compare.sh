#!/bin/bash
# GPU box: the reference's own CUDA build (oracle/_ref/parallel, block 16) against build/parallel on
# the reference's test inputs (tests/params/c{1..}.in = values of the reference's tests/{1..}.in).
# Both print U-CENTER / V-CENTER on stdout and the accumulated SOR seconds on stderr.
cd "$(dirname "$0")/.."
for k in ${@:-1 2}; do
  f=tests/params/c$k.in
  echo "== $f"
  if [ -x oracle/_ref/parallel ]; then
    s=$(date +%s.%N); timeout 900 oracle/_ref/parallel $f 16 > gpurun_out/refcuda_$k.out 2> gpurun_out/refcuda_$k.err; e=$(date +%s.%N)
    echo "reference CUDA (sm_100, block 16): $(tr '\n' ' ' < gpurun_out/refcuda_$k.out) SOR seconds $(tail -c 12 gpurun_out/refcuda_$k.err) wall $(echo "$e - $s" | bc)"
  fi
  s=$(date +%s.%N); timeout 900 build/parallel $f 16 > gpurun_out/ours_$k.out 2> gpurun_out/ours_$k.err; e=$(date +%s.%N)
  echo "nsb200 build/parallel           : $(tr '\n' ' ' < gpurun_out/ours_$k.out) SOR seconds $(tail -c 12 gpurun_out/ours_$k.err) wall $(echo "$e - $s" | bc)"
done
