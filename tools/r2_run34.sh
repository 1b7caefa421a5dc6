set -x
timeout 300 python tools/sanitize_driver.py > gpurun_out/r2_s34_tour_plain.log 2>&1; echo "plain rc=$?"; tail -5 gpurun_out/r2_s34_tour_plain.log
