timeout 900 python -m pytest tests/test_generator_gpu.py -q -x > gpurun_out/t_gc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_gc.log
tail -n 3 gpurun_out/t_gc.log
timeout 300 python profiles/run_gen.py llama-3.2-1b 64 256 32 5 > gpurun_out/gen64_gc.log 2>&1
timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 32 4 > gpurun_out/genq64_gc.log 2>&1
timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 32 3 > gpurun_out/gen512_gc.log 2>&1
tail -n 4 gpurun_out/gen64_gc.log gpurun_out/genq64_gc.log; tail -n 2 gpurun_out/gen512_gc.log
