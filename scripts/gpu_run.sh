timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_generator_gpu.py -q -x > gpurun_out/t_tr.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tr.log
tail -n 6 gpurun_out/t_tr.log
timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 8 3 > gpurun_out/gen512_tr.log 2>&1
timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 8 3 > gpurun_out/genq_tr.log 2>&1
timeout 300 python profiles/run_gen.py llama-3.2-1b 64 256 8 3 > gpurun_out/gen64_tr.log 2>&1
tail -n 2 gpurun_out/gen512_tr.log gpurun_out/genq_tr.log gpurun_out/gen64_tr.log
