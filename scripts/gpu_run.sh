timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_generator_gpu.py -q -x > gpurun_out/t_sk.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_sk.log
tail -n 4 gpurun_out/t_sk.log
timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 32 3 > gpurun_out/gen512_sk.log 2>&1
timeout 300 python profiles/run_gen.py llama-3.2-1b 300 256 32 3 > gpurun_out/gen300_sk.log 2>&1
tail -n 3 gpurun_out/gen512_sk.log gpurun_out/gen300_sk.log
