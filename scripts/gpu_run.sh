timeout 900 python -m pytest tests/test_generator_gpu.py tests/test_attn_gpu.py -q -x > gpurun_out/t_ap.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_ap.log
tail -n 4 gpurun_out/t_ap.log
timeout 300 python profiles/run_gen.py llama-3.2-1b 512 256 32 3 > gpurun_out/gen512_ap.log 2>&1
timeout 300 python profiles/run_gen.py qwen2.5-0.5b 64 256 8 3 > gpurun_out/genq_ap.log 2>&1
tail -n 2 gpurun_out/gen512_ap.log gpurun_out/genq_ap.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"attn_prefill|rmsnorm_row" -c 6 --csv --log-file gpurun_out/l2.csv python profiles/run_gen.py llama-3.2-1b 512 256 2 1 > /dev/null 2>&1
grep -v "^==" gpurun_out/l2.csv | python -c "
import csv,sys
rows=[r for r in csv.reader(sys.stdin) if len(r)>10]
h=rows[0]; k=h.index('Kernel Name'); v=h.index('Metric Value')
for r in rows[1:]: print(r[k][:50], r[v])
"
