#!/bin/bash
# 8-GPU box, short form: multi-GPU parity tests, then weak scaling (8192^2 cells per GPU) at N = 4 and 8
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_8c.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/pytest_multi_8c.log
for N in 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/r1h_scale_n$N.json 2> gpurun_out/r1h_scale_n$N.err; echo "n$N rc=$?"
done
python - <<'PY'
import json
for f in ("r1h_scale_n4","r1h_scale_n8"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, "value %.0f" % d["value"], "ms/step %.2f" % d["ms_per_step"], "frac %.3f" % d["roofline"]["frac"], "e2e", d["e2e"] and "%.0f" % d["e2e"]["value"], d["clocks"]["reasons"])
    except Exception as e:
        print(f, "failed", e)
PY
