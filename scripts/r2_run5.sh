cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_generator_gpu.py tests/test_gemm_gpu.py tests/test_parity_full_gpu.py -x -q > gpurun_out/r2_pytest5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest5.log
grep -E "passed|failed|rc=|Error|error|assert" gpurun_out/r2_pytest5.log | tail -n 15 | cut -c1-300
KX_DEBUG=no-prefetch timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 1 > gpurun_out/r2_trace_b64_async_nopf.txt 2>&1; tail -n 25 gpurun_out/r2_trace_b64_async_nopf.txt | cut -c1-220
timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 1 > gpurun_out/r2_trace_b64_async.txt 2>&1; tail -n 9 gpurun_out/r2_trace_b64_async.txt | cut -c1-220
for b in 1 64 128; do KX_DEBUG=no-prefetch timeout 300 python profiles/run_gen.py llama-3.2-1b $b 256 32 3 2>&1 | tail -n 1 | cut -c1-250; done
for b in 64; do timeout 300 python profiles/run_gen.py llama-3.2-1b $b 256 32 3 2>&1 | tail -n 1 | cut -c1-250; done
