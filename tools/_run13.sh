mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fp8_gpu.py tests/test_mlp_stage.py tests/test_kernels_gpu.py -m gpu -q --timeout 300 > gpurun_out/r2_fp8_tests_n1.log 2>&1; echo rc=$?; tail -4 gpurun_out/r2_fp8_tests_n1.log
python tools/fp8_train_bench.py > gpurun_out/r2_fp8_train_tma.log 2>&1; tail -3 gpurun_out/r2_fp8_train_tma.log
