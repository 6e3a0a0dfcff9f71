timeout 600 python -m pytest tests/test_generator_gpu.py -q -x > gpurun_out/t_tc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tc.log
tail -n 3 gpurun_out/t_tc.log | cut -c1-220
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:attn_prefill -c 2 --csv --log-file gpurun_out/l3.csv python profiles/run_gen.py llama-3.2-1b 128 256 2 1 > /dev/null 2>&1
grep -v "^==" gpurun_out/l3.csv | tail -n 2 | cut -d, -f5,15 | cut -c1-120
