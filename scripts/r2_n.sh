cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2h_n$N.json 2> gpurun_out/r2h_n$N.err
echo "n$N rc=$?"; cut -c1-200 gpurun_out/r2h_n$N.json; tail -n 2 gpurun_out/r2h_n$N.err | cut -c1-200
