set -x
timeout 600 python bench.py > gpurun_out/bench_e_default.json 2> gpurun_out/bench_e_default.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_e_default.json | cut -c1-600
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
timeout 300 $B > gpurun_out/plain_e.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_e.csv $B > gpurun_out/ncu_launch_e.log 2>&1
timeout 600 python bench.py --workload k4 --steps 10 --warmup 3 --no-e2e --no-stream --no-cpu-baseline > gpurun_out/bench_e_k4.json 2> gpurun_out/bench_e_k4.err; tail -1 gpurun_out/bench_e_k4.json | cut -c1-300
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_e_reference.json 2>&1; tail -1 gpurun_out/bench_e_reference.json | cut -c1-400
