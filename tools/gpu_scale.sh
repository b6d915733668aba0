#!/bin/bash
# 8-GPU box: multi-GPU parity tests, then weak scaling at N = 1, 2, 4, 8 and strong scaling 16384^2 at 8
nvidia-smi topo -m > gpurun_out/topo_8.txt 2>&1
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_8.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_multi_8.log
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "n$N rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 3 --warmup 3 --scaling strong --size 16384 --no-e2e > gpurun_out/strong16k_n8.json 2> gpurun_out/strong16k_n8.err; echo "strong rc=$?"
python - <<'PY'
import json
for f in ("scale_n1","scale_n2","scale_n4","scale_n8","strong16k_n8"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, "value %.0f" % d["value"], "ms/step %.2f" % d["ms_per_step"], "frac %.3f" % d["roofline"]["frac"], "e2e", d["e2e"] and "%.0f" % d["e2e"]["value"], d["clocks"]["reasons"])
    except Exception as e:
        print(f, "failed", e)
PY
