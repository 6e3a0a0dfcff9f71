timeout 900 python -m pytest tests/test_strings_gpu.py tests/test_zz_no_alloc.py tests/test_pipeline.py -q -x -m gpu > gpurun_out/t_k9.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_k9.log
tail -n 5 gpurun_out/t_k9.log
timeout 600 python bench.py --workload validate_hash_10m > gpurun_out/bench_k9_v5.json 2> gpurun_out/bench_k9_v5.err; echo "bench rc=$?"
cut -c1-700 gpurun_out/bench_k9_v5.json
