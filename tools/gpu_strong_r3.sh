#!/bin/bash
# 8-GPU box: multi-GPU parity tests with the planned band heights, strong scaling of 16384^2 at N = 8 (and 4), weak N = 8.
NG=$(nvidia-smi -L | wc -l)
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r3q_pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r3q_pytest_multi.log
for N in 8 4; do
  [ $N -le $NG ] || continue
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --steps 5 --warmup 3 --size 16384 --scaling strong --no-e2e > gpurun_out/r3q_strong_n$N.json 2> gpurun_out/r3q_strong_n$N.err; echo "strong n$N rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29559 bench.py --gpus $NG --steps 5 --warmup 3 --no-e2e > gpurun_out/r3q_weak_n$NG.json 2> gpurun_out/r3q_weak_n$NG.err; echo "weak rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3q_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        r=d["roofline"]
        print(f, "value %.0f" % d["value"], "ms/step %.2f" % d["ms_per_step"], "frac %.3f" % r["frac"], "k", r.get("iterations_per_pass"), "chunk", r.get("chunk_rows"), "us/it %.1f" % (1e3 * r["ms_per_iteration"]), "parity", d.get("parity", {}).get("ok"), d["clocks"]["reasons"])
    except Exception as e:
        print(f, "failed", e)
PY
