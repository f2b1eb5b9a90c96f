B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
for tb in 40 60; do for w in k2 k4; do
echo "== f32 TB=$tb $w"; HGA_B200_PIPE_TB=$tb timeout 300 $B --workload $w --precision f32 2>gpurun_out/err_u.txt | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['other_precisions'])" || tail -5 gpurun_out/err_u.txt
done; done
HGA_B200_PIPE_TB=60 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
HGA_B200_PIPE_TB=80 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
