./tools/ubench_ffma2.bin > gpurun_out/r2_s51_ubench_ffma2.txt 2>&1; cat gpurun_out/r2_s51_ubench_ffma2.txt
