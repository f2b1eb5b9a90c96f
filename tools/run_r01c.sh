set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/gpu_tests_c.log
cat gpurun_out/gpu_tests_c.log
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
for nf in 4 8; do
  for prec in f64 f32; do
    echo "== NF=$nf $prec"
    HGA_B200_PIPE_NF=$nf timeout 300 $B --precision $prec 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d.get('other_precisions'))"
  done
done
HGA_B200_PIPE_NF=4 HGA_B200_PROF=1 timeout 300 $B --precision f64 > /dev/null 2> gpurun_out/prof_c_f64_nf4.txt
HGA_B200_PIPE_NF=8 HGA_B200_PROF=1 timeout 300 $B --precision f32 > /dev/null 2> gpurun_out/prof_c_f32_nf8.txt
tail -13 gpurun_out/prof_c_f64_nf4.txt; tail -13 gpurun_out/prof_c_f32_nf8.txt
