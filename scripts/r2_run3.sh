cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest3.log
grep -E "passed|failed|rc=|Error|error|assert" gpurun_out/r2_pytest3.log | tail -n 15 | cut -c1-300
timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 2 > gpurun_out/r2_trace_b64_push.txt 2>&1; tail -n 34 gpurun_out/r2_trace_b64_push.txt | cut -c1-200
KX_DEBUG=no-prefetch timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 2 > gpurun_out/r2_trace_b64_push_nopf.txt 2>&1; tail -n 34 gpurun_out/r2_trace_b64_push_nopf.txt | cut -c1-200
for b in 1 64 128; do KX_DEBUG=no-prefetch timeout 300 python profiles/run_gen.py llama-3.2-1b $b 256 32 3 2>&1 | tail -n 1 | cut -c1-250; done
for b in 1 64 128; do timeout 300 python profiles/run_gen.py llama-3.2-1b $b 256 32 3 2>&1 | tail -n 1 | cut -c1-250; done
