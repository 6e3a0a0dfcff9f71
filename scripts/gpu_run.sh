timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/t_gpu_all.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_gpu_all.log
tail -n 5 gpurun_out/t_gpu_all.log | cut -c1-200
KX_ATTN_TC=0 timeout 600 python -m pytest tests/test_generator_gpu.py -q -x -k "not tcgen05" > gpurun_out/t_notc.log 2>&1; tail -n 2 gpurun_out/t_notc.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
