export NSB_SPIN_TIMEOUT_MS=600000
for tool in memcheck racecheck initcheck synccheck; do
  echo "==== compute-sanitizer --tool $tool python tools/sanitize_case.py" > gpurun_out/r2h_sanitizer_$tool.txt
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py >> gpurun_out/r2h_sanitizer_$tool.txt 2>&1
  echo "exit code $?" >> gpurun_out/r2h_sanitizer_$tool.txt
  tail -4 gpurun_out/r2h_sanitizer_$tool.txt
done
