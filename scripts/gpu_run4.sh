timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
cut -c1-220 gpurun_out/bench_n8.json; tail -n 2 gpurun_out/bench_n8.err | cut -c1-200
