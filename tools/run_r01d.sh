set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/gpu_tests_d.log
cat gpurun_out/gpu_tests_d.log
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
timeout 300 $B 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d.get('other_precisions'))"
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
timeout 300 $B > gpurun_out/plain_d.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_d_f64 $B > gpurun_out/ncu_d_f64.log 2>&1
tail -2 gpurun_out/ncu_d_f64.log
timeout 300 $B --precision f32 > gpurun_out/plain_d32.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_d_f32 $B --precision f32 > gpurun_out/ncu_d_f32.log 2>&1
tail -2 gpurun_out/ncu_d_f32.log
