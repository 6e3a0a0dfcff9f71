timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_generator_gpu.py -q -x > gpurun_out/t_am.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_am.log
tail -n 3 gpurun_out/t_am.log
for f in 0 1; do
KX_FUSED_ARGMAX=$f timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 32 3 > gpurun_out/genq64_f$f.log 2>&1
KX_FUSED_ARGMAX=$f timeout 300 python profiles/run_gen.py llama-3.2-1b 64 256 32 4 > gpurun_out/gen64_f$f.log 2>&1
KX_FUSED_ARGMAX=$f timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 32 3 > gpurun_out/gen512_f$f.log 2>&1
done
tail -n 2 gpurun_out/genq64_f*.log gpurun_out/gen64_f*.log gpurun_out/gen512_f*.log
