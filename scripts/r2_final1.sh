cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest_final.log 2>&1 ) 2>&1 | grep real
echo "pytest rc=$?"; tail -n 4 gpurun_out/r2_pytest_final.log
b() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/r2f_$name.json 2> gpurun_out/r2f_$name.err; echo "$name rc=$?"; cut -c1-260 gpurun_out/r2f_$name.json; }
b n1 --steps 5 --warmup 3
b b64 --requests 64 --steps 5 --warmup 3
b qwen --arch qwen2.5-0.5b --requests 64 --seed 3 --steps 5 --warmup 3
b k9 --workload validate_hash_10m
b ref --impl reference --steps 2 --warmup 1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
for B in 64 512; do
timeout 300 python profiles/run_gen.py llama-3.2-1b $B 256 4 1 > gpurun_out/plain_b$B.log 2>&1 &&
timeout 900 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/r2f_launches_b$B.csv python profiles/run_gen.py llama-3.2-1b $B 256 4 1 > gpurun_out/ncu_b$B.log 2>&1
echo "ncu B=$B rc=$?"
done
