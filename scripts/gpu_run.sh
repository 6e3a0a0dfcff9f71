timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/t_gpu_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_gpu_all.log
tail -n 5 gpurun_out/t_gpu_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/bench_default.json
timeout 900 python bench.py --requests 64 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err
timeout 900 python bench.py --arch qwen2.5-0.5b --requests 64 > gpurun_out/bench_qwen.json 2> gpurun_out/bench_qwen.err
cut -c1-300 gpurun_out/bench_b64.json; cut -c1-300 gpurun_out/bench_qwen.json
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4500 --csv --log-file gpurun_out/bench_launches.csv python bench.py --steps 1 --warmup 1 --cpu-requests 1 > gpurun_out/bench_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/bench_launches.csv
