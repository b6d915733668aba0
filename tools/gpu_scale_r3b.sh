#!/bin/bash
# 8-GPU box, final tree: multi-GPU parity tests, weak scaling (8192^2 cells per GPU) at N = 1, 2, 4, 8, strong 16384^2 at N = 8.
NG=$(nvidia-smi -L | wc -l)
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r3x_pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r3x_pytest_multi.log
timeout 300 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r3x_scale_n1.json 2> gpurun_out/r3x_scale_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  [ $N -le $NG ] || continue
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r3x_scale_n$N.json 2> gpurun_out/r3x_scale_n$N.err; echo "n$N rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29569 bench.py --gpus $NG --steps 5 --warmup 3 --size 16384 --scaling strong --no-e2e > gpurun_out/r3x_strong_n$NG.json 2> gpurun_out/r3x_strong_n$NG.err; echo "strong rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3x_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        r=d["roofline"]
        print(f, "value %.0f" % d["value"], "ms/step %.2f" % d["ms_per_step"], "frac %.3f" % r["frac"], "k", r.get("iterations_per_pass"), "chunk", r.get("chunk_rows"), "us/it %.1f" % (1e3 * r["ms_per_iteration"]), "e2e", d.get("e2e") and "%.0f" % d["e2e"]["value"], "parity", d.get("parity", {}).get("ok"), d["clocks"]["reasons"])
    except Exception as e:
        print(f, "failed", e)
PY
