#!/bin/bash
# usage: tools/chunk_sweep.sh "<variants>" "<chunk rows list>" [bench args]
cd "$(dirname "$0")/.."
for v in $1; do for ch in $2; do
  NSB200_LIB=$PWD/build/variants/lib_$v.so python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --chunk-rows $ch "${@:3}" > /tmp/cs.json 2>/tmp/cs.err
  python - "$v" "$ch" <<'PY'
import json,sys
try:
    d=json.load(open("/tmp/cs.json")); print("%s chunk %s: value %.0f ms/step %.2f kernel %.1f us frac %.3f" % (sys.argv[1], sys.argv[2], d["value"], d["ms_per_step"], d["roofline"]["launch_ms"]*1e3, d["roofline"]["frac"]))
except Exception as e: print(sys.argv[1], sys.argv[2], "failed", e, open("/tmp/cs.err").read()[-300:])
PY
done; done
