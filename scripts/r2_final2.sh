cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_tp_gpu.py -x -q -s > gpurun_out/r2_tp_tests_v3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tp_tests_v3.log
grep -E "passed|failed|rc=|llama-3.1-8b|Error|error|assert" gpurun_out/r2_tp_tests_v3.log | tail -n 8 | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2f_n2.json 2> gpurun_out/r2f_n2.err
echo "n2 rc=$?"; cut -c1-200 gpurun_out/r2f_n2.json; tail -n 2 gpurun_out/r2f_n2.err | cut -c1-200
