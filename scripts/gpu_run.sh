timeout 600 python -m pytest tests/test_generator_gpu.py -q -x > gpurun_out/t_tc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tc.log
tail -n 6 gpurun_out/t_tc.log | cut -c1-220
KX_ATTN_TC=1 timeout 600 python -m pytest tests/test_generator_gpu.py -q -x -k "qwen25 or config3 or wide or chunked" > gpurun_out/t_tc2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tc2.log
tail -n 3 gpurun_out/t_tc2.log | cut -c1-220
