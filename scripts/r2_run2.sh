cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_pytest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest2.log
grep -E "passed|failed|rc=|Error|error|assert" gpurun_out/r2_pytest2.log | tail -n 15 | cut -c1-300
grep -E "^(llama|qwen|tiny)" gpurun_out/r2_pytest2.log | cut -c1-250
timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 2 > gpurun_out/r2_trace_b64_fused.txt 2>&1; tail -n 34 gpurun_out/r2_trace_b64_fused.txt | cut -c1-200
KX_DEBUG=no-prefetch timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 1 > gpurun_out/r2_trace_b64_fused_nopf.txt 2>&1; tail -n 9 gpurun_out/r2_trace_b64_fused_nopf.txt | cut -c1-200
timeout 300 python profiles/trace_decode.py llama-3.2-1b 8 256 8 1 > gpurun_out/r2_trace_b8_fused.txt 2>&1; tail -n 9 gpurun_out/r2_trace_b8_fused.txt | cut -c1-200
timeout 300 python profiles/trace_decode.py qwen2.5-0.5b 64 256 8 1 > gpurun_out/r2_trace_qwen_b64_fused.txt 2>&1; tail -n 9 gpurun_out/r2_trace_qwen_b64_fused.txt | cut -c1-200
for b in 1 64 128; do timeout 300 python profiles/run_gen.py llama-3.2-1b $b 256 32 3 2>&1 | tail -n 2 | cut -c1-250; done
