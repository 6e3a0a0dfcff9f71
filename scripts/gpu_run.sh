timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_gpu_all.log
tail -n 4 gpurun_out/t_gpu_all.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_ref.json
