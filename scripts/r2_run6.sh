cd $GRAFT_REPO_ROOT
KX_DEBUG=no-prefetch timeout 300 python profiles/trace_decode.py llama-3.2-1b 64 256 8 1 > gpurun_out/r2_trace_b64_probe.txt 2>&1; tail -n 25 gpurun_out/r2_trace_b64_probe.txt | cut -c1-220
