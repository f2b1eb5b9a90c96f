timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
for cfg in "2 4 20 9" "2 4 40 9"; do set -- $cfg; echo "ver=$1 S=$2 TB=$3 OCC=$4"; HGA_B200_PIPE_V=$1 HGA_B200_PIPE_S=$2 HGA_B200_PIPE_TB=$3 HGA_B200_PIPE_OCC=$4 timeout 300 $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); o=d['other_precisions']['f32']; print('f64', d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['warmup_samples'], '| f32', o['value']/1e9, o['ms_per_step'], o['roofline_frac'], o['max_abs_diff_vs_headline'])"; done > gpurun_out/sweep_v2.txt 2>&1
cat gpurun_out/sweep_v2.txt
for p in f64 f32; do echo "== default shapes, headline $p (last lines = the OTHER precision)"; HGA_B200_PROF=1 timeout 300 $B --precision $p 2>&1 | grep "hga prof" | tail -9; done > gpurun_out/prof2.txt
cat gpurun_out/prof2.txt
