set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gpu_tests_b.log
cat gpurun_out/gpu_tests_b.log
timeout 600 python bench.py > gpurun_out/bench_b_default.json 2> gpurun_out/bench_b_default.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_b_default.json
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
HGA_B200_FORM=64 timeout 300 $B 2>/dev/null | tail -1 | cut -c1-200
HGA_B200_PROF=1 timeout 300 $B > /dev/null 2> gpurun_out/prof_b.txt; tail -20 gpurun_out/prof_b.txt
