#!/bin/bash
# SASS evidence per kernel of libnsb200.so (no GPU needed): instruction counts that show how the hot kernels move data.
#   UBLKCP = cp.async.bulk (TMA engine, global -> shared), SYNCS = mbarrier operations, ELECT = elect.sync,
#   UCGABAR/CGABAR = cluster barriers, STG.E.128 / LDG.E.128 = 16-byte global accesses, DADD/DMUL/DFMA = FP64 pipe.
# usage: tools/sass_summary.sh > profiles/sass_summary.txt
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SO=$ROOT/navierstokes-parallel_b200/libnsb200.so
echo "cuobjdump -sass $SO ($(cuobjdump -lelf $SO | grep -c sm_100a) sm_100a cubins; nvcc $(nvcc --version | grep -o 'release [0-9.]*'))"
echo "kernel | instructions | UBLKCP | SYNCS | ELECT | U?CGABAR | LDS | LDG.E.128 | LDG(other) | STG.E.128 | STG(other) | SHFL | DADD+DMUL+DFMA | MUFU.RCP64H"
cuobjdump -sass "$SO" | awk '
  /Function :/ { if (name != "") flush(); name=$3; n=0; ub=0; sy=0; el=0; cg=0; lds=0; l128=0; lo=0; s128=0; so=0; sh=0; fp=0; mu=0; next }
  /^[ \t]+\/\*[0-9a-f]+\*\// {
     n++
     if ($0 ~ /UBLKCP/) ub++
     if ($0 ~ /SYNCS/) sy++
     if ($0 ~ / ELECT /) el++
     if ($0 ~ /CGABAR/) cg++
     if ($0 ~ / LDS/) lds++
     if ($0 ~ / LDG\.E\.128| LDG\.E\.[A-Z.]*128/) l128++; else if ($0 ~ / LDG/) lo++
     if ($0 ~ / STG\.E\.128| STG\.E\.[A-Z.]*128/) s128++; else if ($0 ~ / STG/) so++
     if ($0 ~ / SHFL/) sh++
     if ($0 ~ / DADD | DMUL | DFMA /) fp++
     if ($0 ~ /MUFU\.RCP64H/) mu++
  }
  function flush() { printf "%s | %d | %d | %d | %d | %d | %d | %d | %d | %d | %d | %d | %d | %d\n", name, n, ub, sy, el, cg, lds, l128, lo, s128, so, sh, fp, mu }
  END { if (name != "") flush() }' | c++filt | sed 's/(SorPersistArgs)//; s/(SorStreamArgs)//' | sort
