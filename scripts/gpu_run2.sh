timeout 900 python -m pytest tests/test_tp_gpu.py -q -x > gpurun_out/t_tp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_tp.log
tail -n 15 gpurun_out/t_tp.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
cut -c1-600 gpurun_out/bench_n2.json; tail -n 3 gpurun_out/bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err
cut -c1-400 gpurun_out/bench_ref_n2.json
