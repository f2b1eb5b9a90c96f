set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/gpu_tests_g.log
cat gpurun_out/gpu_tests_g.log
timeout 600 python bench.py > gpurun_out/bench_g_default.json 2> gpurun_out/bench_g_default.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_g_default.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['roofline']['frac'], json.dumps(d['stream_latency']))"
tail -3 gpurun_out/bench_g_default.err
