set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gpu_tests.log
cat gpurun_out/gpu_tests.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_default.json
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
timeout 300 $B > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_r01.csv $B > gpurun_out/ncu_launch.log 2>&1
timeout 300 $B > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_pipe_f64 $B > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
timeout 300 $B --precision f32 > gpurun_out/plain_f32.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_pipe_f32 $B --precision f32 > gpurun_out/ncu_f32.log 2>&1
tail -2 gpurun_out/ncu_f32.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2>&1; tail -1 gpurun_out/bench_reference.json
