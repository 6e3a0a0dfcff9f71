SPLITS=1,2,4,8,12 timeout 300 python profiles/sweep_decode_gemm.py 1b.qkv 1b.wo 1b.wgu 1b.wd q.wgu q.wd > gpurun_out/sweep_ring72.log 2>&1
cat gpurun_out/sweep_ring72.log
timeout 300 python profiles/run_gen.py llama-3.2-1b 64 256 32 4 > gpurun_out/gen64_ring72.log 2>&1
timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 32 3 > gpurun_out/genq64_ring72.log 2>&1
tail -n 2 gpurun_out/gen64_ring72.log gpurun_out/genq64_ring72.log
