timeout 900 python -m pytest tests/test_generator_gpu.py tests/test_pipeline.py -q -x -m gpu > gpurun_out/t_tc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tc.log
tail -n 4 gpurun_out/t_tc.log | cut -c1-220
