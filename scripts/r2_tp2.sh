cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests/test_tp_gpu.py -x -q -s > gpurun_out/r2_tp_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tp_tests.log
grep -E "passed|failed|rc=|llama-3.1-8b|Error|error" gpurun_out/r2_tp_tests.log | tail -n 12 | cut -c1-300
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --arch llama-3.1-8b --tp 2 --prefix-len 1024 --requests 64 --prompt-len 32 --new-tokens 64 --steps 3 --warmup 3 --cpu-requests 2 --seed 5 > gpurun_out/r2_bench_cfg5_2gpu.json 2> gpurun_out/r2_bench_cfg5_2gpu.err ) 2>&1 | grep real
cut -c1-300 gpurun_out/r2_bench_cfg5_2gpu.json; tail -n 3 gpurun_out/r2_bench_cfg5_2gpu.err | cut -c1-300
