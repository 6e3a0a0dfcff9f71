timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/t_gpu_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_gpu_all.log
tail -n 5 gpurun_out/t_gpu_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/bench_default.json
timeout 900 python bench.py --requests 64 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err
timeout 900 python bench.py --arch qwen2.5-0.5b --requests 64 > gpurun_out/bench_qwen.json 2> gpurun_out/bench_qwen.err
cut -c1-300 gpurun_out/bench_b64.json; cut -c1-300 gpurun_out/bench_qwen.json
