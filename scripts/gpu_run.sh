for b in 1 8 32; do timeout 300 python profiles/run_gen.py llama-3.2-1b $b 256 32 3 2>&1 | tail -n 1; done
for b in 1 8; do timeout 300 python profiles/run_gen.py qwen2.5-0.5b $b 256 32 3 2>&1 | tail -n 1; done
