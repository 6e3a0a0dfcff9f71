timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/t_gpu_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_gpu_all.log
tail -n 6 gpurun_out/t_gpu_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 4 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cat gpurun_out/bench_default.json | cut -c1-1500
tail -n 3 gpurun_out/bench_default.err
