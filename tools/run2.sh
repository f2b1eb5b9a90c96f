set -x
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline"
for cfg in "1 4 40 9" "1 4 40 1" "1 4 20 9" "1 2 40 9" "1 8 40 9"; do set -- $cfg; echo "variant=$1 S=$2 TB=$3 OCC=$4"; HGA_B200_VARIANT=$1 HGA_B200_PIPE_S=$2 HGA_B200_PIPE_TB=$3 HGA_B200_PIPE_OCC=$4 $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'])"; done > gpurun_out/sweep2.txt 2>&1
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r01.csv $B > gpurun_out/ncu_launch.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe -s 3 -c 1 -o gpurun_out/prof_pipe_f64 $B > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
cat gpurun_out/sweep2.txt
