timeout 900 python -m pytest tests/test_tp_gpu.py -q -x > gpurun_out/t_tp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tp.log
tail -n 5 gpurun_out/t_tp.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --arch llama-3.1-8b --tp 2 --prefix-len 1024 --requests 64 --prompt-len 32 --new-tokens 64 --steps 3 --warmup 3 > gpurun_out/bench_cfg5_v3.json 2> gpurun_out/bench_cfg5_v3.err
tail -n 3 gpurun_out/bench_cfg5_v3.json; tail -n 5 gpurun_out/bench_cfg5_v3.err
