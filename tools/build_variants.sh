#!/bin/bash
# Builds tuning variants of libnsb200.so into build/variants/ (they travel to the GPU box with
# gpurun; bench.py / tests load one with NSB200_LIB=<path>).  usage: tools/build_variants.sh name "flags" ...
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$ROOT/build/variants"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  make -s -C "$ROOT/navierstokes-parallel_b200/csrc" lib OUT="$ROOT/build/variants/lib_$name.so" \
       OBJDIR="$ROOT/build/obj_$name" VARIANT_FLAGS="$flags"
  echo "== $name ($flags)"; grep -h "registers\|spill" "$ROOT/build/obj_$name/sor_stream.ptxas.log" | head -2
done
