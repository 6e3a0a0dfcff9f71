timeout 600 python -m pytest tests/test_generator_gpu.py -q -x > gpurun_out/t_tc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tc.log
tail -n 3 gpurun_out/t_tc.log | cut -c1-220
timeout 300 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:attn_prefill -c 2 --csv --log-file gpurun_out/l3.csv python profiles/run_gen.py llama-3.2-1b 128 256 2 1 > /dev/null 2>&1
grep -v "^==" gpurun_out/l3.csv | python -c "
import csv,sys
rows=[r for r in csv.reader(sys.stdin) if len(r)>10]
h=rows[0]; k=h.index('Kernel Name'); v=h.index('Metric Value'); m=h.index('Metric Name'); g=h.index('Grid Size')
for r in rows[1:]: print(r[k][:40], r[g], r[m][:30], r[v])
"
