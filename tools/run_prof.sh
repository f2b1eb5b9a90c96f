B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
tools/microbench/systolic_rates > gpurun_out/systolic_rates.txt 2>&1
for p in f64 f32; do HGA_B200_PROF=1 $B --precision $p 2>&1 | grep "hga prof" | tail -8 > gpurun_out/prof_$p.txt; done
cat gpurun_out/prof_f64.txt gpurun_out/prof_f32.txt
for cfg in "f64 4 40" "f64 2 40" "f64 4 20" "f32 4 40" "f32 2 40" "f32 4 20"; do set -- $cfg; echo "prec=$1 S=$2 TB=$3"; HGA_B200_PIPE_S=$2 HGA_B200_PIPE_TB=$3 $B --precision $1 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['warmup_samples'])"; done > gpurun_out/sweep.txt 2>&1
cat gpurun_out/sweep.txt
