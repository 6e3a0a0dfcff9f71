set -x
timeout 900 python -m pytest tests/test_attn_gpu.py -q > gpurun_out/t_attn.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_attn.log
tail -n 40 gpurun_out/t_attn.log
