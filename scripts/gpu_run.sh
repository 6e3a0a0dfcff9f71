timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cut -c1-260 gpurun_out/bench_default.json
timeout 900 python bench.py --requests 64 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err
timeout 900 python bench.py --arch qwen2.5-0.5b --requests 64 > gpurun_out/bench_qwen.json 2> gpurun_out/bench_qwen.err
timeout 600 python bench.py --workload validate_hash_10m > gpurun_out/bench_k9.json 2> gpurun_out/bench_k9.err
cut -c1-260 gpurun_out/bench_b64.json; cut -c1-260 gpurun_out/bench_qwen.json; cut -c1-200 gpurun_out/bench_k9.json
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4500 --csv --log-file gpurun_out/bench_launches.csv python bench.py --steps 1 --warmup 1 --cpu-requests 1 > gpurun_out/bench_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/bench_launches.csv
