mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stress_gpu.py tests/test_multigpu.py -m gpu -q --timeout 300 -k "stress or stalled or wrap or delays" > gpurun_out/r2_stress_n2.log 2>&1; echo rc=$?; tail -15 gpurun_out/r2_stress_n2.log
timeout 900 python -m pytest tests/test_multigpu.py tests/test_tp_gpu.py tests/test_p2p_native_gpu.py tests/test_native_exec.py -m gpu -q --timeout 400 -k "not stalled" > gpurun_out/r2_gpu_tests_n2.log 2>&1; echo rc=$?; tail -8 gpurun_out/r2_gpu_tests_n2.log
