mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fp8_gpu.py tests/test_mlp_stage.py -m gpu -q --timeout 300 > gpurun_out/r2_fp8_tests2_n1.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2_fp8_tests2_n1.log
for red in 0 1; do echo "SDML_W_GROUP=2 SDML_FP8_TMA_REDUCE=$red"; SDML_W_GROUP=2 SDML_FP8_TMA_REDUCE=$red python tools/fp8_train_bench.py 2>&1 | tail -3; done | tee gpurun_out/r2_fp8_wgroup2_ab.log
