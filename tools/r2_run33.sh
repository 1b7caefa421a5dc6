set -x
timeout 300 python tools/sanitize_driver.py > gpurun_out/r2_s33_tour_plain.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/r2_s33_tour_plain.log
for tool in memcheck racecheck initcheck synccheck; do
timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 python tools/sanitize_driver.py > gpurun_out/r2_s33_sanitizer_$tool.log 2>&1; echo "$tool rc=$?"; tail -4 gpurun_out/r2_s33_sanitizer_$tool.log
done
