set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,memory.total --format=csv | tail -1
nproc; free -g | head -2
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest1.log
tail -n 15 gpurun_out/r2_pytest1.log | cut -c1-300
timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 2 > gpurun_out/r2_trace_b64_base.txt 2>&1; tail -n 45 gpurun_out/r2_trace_b64_base.txt | cut -c1-200
timeout 300 python profiles/trace_decode.py llama-3.2-1b 512 256 8 1 > gpurun_out/r2_trace_b512_base.txt 2>&1; tail -n 14 gpurun_out/r2_trace_b512_base.txt | cut -c1-200
timeout 300 python profiles/l2_prefetch_bench.py 64 > gpurun_out/r2_l2_prefetch_b64.txt 2>&1; cat gpurun_out/r2_l2_prefetch_b64.txt | cut -c1-200
