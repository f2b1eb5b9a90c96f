set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/gpu_tests_x.log
cat gpurun_out/gpu_tests_x.log
timeout 600 python bench.py > gpurun_out/bench_x_default.json 2> gpurun_out/bench_x_default.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_x_default.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['other_precisions'], d['e2e']['value'])"
timeout 600 python bench.py --workload k4 --steps 10 --warmup 3 --no-e2e --no-stream --no-cpu-baseline > gpurun_out/bench_x_k4.json 2>/dev/null; tail -1 gpurun_out/bench_x_k4.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['other_precisions'])"
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
timeout 300 $B > gpurun_out/plain_x.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_x.csv $B > gpurun_out/ncu_launch_x.log 2>&1
timeout 300 $B > gpurun_out/plain_x2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_x_f64 $B > gpurun_out/ncu_x_f64.log 2>&1
tail -2 gpurun_out/ncu_x_f64.log
HGA_B200_PROF=1 timeout 300 $B > /dev/null 2> gpurun_out/prof_x.txt
