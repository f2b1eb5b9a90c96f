# One GPU box, one pass: parity tests, the default bench line, launch list, full ncu captures, auxiliary kernels.
# usage: gpurun --timeout 2400 -- 'bash tools/run_r01_full.sh'      (outputs under gpurun_out/, copy what is judged to profiles/)
set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gpu_tests.log
cat gpurun_out/gpu_tests.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_default.json | cut -c1-400
timeout 600 python bench.py --workload k4 --steps 10 --warmup 3 --no-e2e --no-stream --no-cpu-baseline > gpurun_out/bench_k4.json 2> gpurun_out/bench_k4.err
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
timeout 300 $B > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launch.log 2>&1
timeout 300 $B > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_pipe_f64 $B > gpurun_out/ncu_full.log 2>&1
timeout 300 $B --precision f32 > gpurun_out/plain_f32.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe2 -s 3 -c 1 -o gpurun_out/prof_pipe_f32 $B --precision f32 > gpurun_out/ncu_f32.log 2>&1
HGA_B200_PROF=1 timeout 300 $B > /dev/null 2> gpurun_out/prof_cycles.txt
timeout 300 python tools/bench_aux.py > gpurun_out/bench_aux.json 2> gpurun_out/bench_aux.err
timeout 300 python tools/stream_breakdown.py > gpurun_out/stream_breakdown.txt 2>&1
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2>&1; tail -1 gpurun_out/bench_reference.json | cut -c1-300
