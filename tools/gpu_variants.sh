for v in m0 m1 m2; do
  echo "=== variant $v"
  export NSB200_LIB=$PWD/build/variants/lib_$v.so
  timeout 200 python tools/k_probe.py 2048,4096,8192 1 0 2
  timeout 100 python tools/k_probe.py 8191 1
  timeout 100 python tools/k_probe.py 8192 2
done 2>&1 | tee gpurun_out/r2m_variants.txt
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv,noheader | tee -a gpurun_out/r2m_variants.txt
