#!/bin/bash
# 8-GPU box: strong scaling of BASELINE configs[3] (16384^2 at 1/2/4/8) and the 32768^2 box at 1 and 8 GPUs;
# multi-GPU parity tests first.
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_8b.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/pytest_multi_8b.log
run() {  # size N tag
  if [ $2 = 1 ]; then timeout 600 python bench.py --steps 3 --warmup 3 --size $1 --no-e2e --no-cpu-baseline > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 296$2$2 bench.py --gpus $2 --steps 3 --warmup 3 --scaling strong --size $1 --no-e2e > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "$3 rc=$?"
}
run 16384 1 strong16k_n1; run 16384 2 strong16k_n2; run 16384 4 strong16k_n4; run 16384 8 strong16k_n8
run 32768 8 strong32k_n8; run 32768 1 strong32k_n1
python - <<'PY'
import json
for f in ("strong16k_n1","strong16k_n2","strong16k_n4","strong16k_n8","strong32k_n1","strong32k_n8"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, "value %.0f" % d["value"], "ms/step %.2f" % d["ms_per_step"], "frac %.3f" % d["roofline"]["frac"], d["clocks"]["reasons"])
    except Exception as e:
        print(f, "failed", e)
PY
