set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/gpu_tests_h.log
cat gpurun_out/gpu_tests_h.log
for m in 1 0; do
HGA_B200_STREAM_MAPPED=$m timeout 600 python bench.py --steps 5 --no-e2e --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps(d['stream_latency']))"
done
