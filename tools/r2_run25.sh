set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_oracle_full_size.py tests/test_gpu_shared_frame.py -m gpu -x -q -k "strip or cfg4 or shared" > gpurun_out/r2_s25_pytest.log 2>&1; tail -4 gpurun_out/r2_s25_pytest.log
