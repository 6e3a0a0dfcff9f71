timeout 900 python -m pytest tests/test_tp_gpu.py -q -x > gpurun_out/t_tp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tp.log
tail -n 4 gpurun_out/t_tp.log | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --arch llama-3.1-8b --tp 2 --prefix-len 1024 --requests 64 --prompt-len 32 --new-tokens 64 --steps 3 --warmup 3 > gpurun_out/bench_cfg5_v4.json 2> gpurun_out/bench_cfg5_v4.err
cut -c1-200 gpurun_out/bench_cfg5_v4.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
cut -c1-200 gpurun_out/bench_n2.json
