for r in 100000000 4096; do
KX_STREAM_OUT_ROWS=$r timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 8 3 > gpurun_out/gen512_so$r.log 2>&1
KX_STREAM_OUT_ROWS=$r timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 8 3 > gpurun_out/genq_so$r.log 2>&1
done
tail -n 2 gpurun_out/gen512_so*.log gpurun_out/genq_so*.log
