timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_generator_gpu.py -q -x > gpurun_out/t_re.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_re.log
tail -n 3 gpurun_out/t_re.log
for w in 0 15; do
KX_PAIR256_MAX_WASTE=$w timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 32 3 > gpurun_out/genq64_w$w.log 2>&1
done
timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 32 3 > gpurun_out/gen512_re.log 2>&1
timeout 300 python profiles/run_gen.py llama-3.2-1b 64 256 32 3 > gpurun_out/gen64_re.log 2>&1
tail -n 2 gpurun_out/genq64_w*.log gpurun_out/gen512_re.log gpurun_out/gen64_re.log
