timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_attn_gpu.py tests/test_generator_gpu.py -q -x > gpurun_out/t_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_all.log
tail -n 6 gpurun_out/t_all.log
SPLITS=8,10,12,16 timeout 300 python profiles/sweep_decode_gemm.py 1b.qkv 1b.wo 1b.wd 8b/2.qkv 8b/2.wo 8b/2.wd q.qkv q.wo q.wd > gpurun_out/sweep16.log 2>&1
cat gpurun_out/sweep16.log
for ms in 8 16; do
KX_MAX_SPLITS=$ms timeout 300 python profiles/run_gen.py llama-3.2-1b 64 256 32 3 > gpurun_out/gen64_s$ms.log 2>&1
KX_MAX_SPLITS=$ms timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 32 3 > gpurun_out/genq64_s$ms.log 2>&1
KX_MAX_SPLITS=$ms timeout 400 python profiles/run_prefix.py llama-3.1-8b 64 1024 32 64 > gpurun_out/prefix_s$ms.log 2>&1
done
timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 32 3 > gpurun_out/gen512.log 2>&1
tail -n 3 gpurun_out/gen64_s*.log gpurun_out/genq64_s*.log gpurun_out/prefix_s*.log gpurun_out/gen512.log
