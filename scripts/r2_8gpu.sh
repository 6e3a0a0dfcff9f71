cd $GRAFT_REPO_ROOT
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --arch llama-3.1-8b --tp 2 --prefix-len 1024 --requests 256 --prompt-len 32 --new-tokens 64 --steps 3 --warmup 3 --cpu-requests 2 --seed 5 > gpurun_out/r2_bench_cfg5_8gpu.json 2> gpurun_out/r2_bench_cfg5_8gpu.err ) 2>&1 | grep real
cut -c1-300 gpurun_out/r2_bench_cfg5_8gpu.json; tail -n 2 gpurun_out/r2_bench_cfg5_8gpu.err | cut -c1-300
