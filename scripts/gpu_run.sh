timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_prefill_tc --launch-skip 2 --launch-count 1 -o gpurun_out/r01_attn_prefill_tc python profiles/run_gen.py llama-3.2-1b 128 256 2 1 > gpurun_out/ncu_tc.log 2>&1
ls -la gpurun_out/r01_attn_prefill_tc.ncu-rep
