B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
for tb in 40 60; do for w in k2 k4; do
echo "== TB=$tb $w"; HGA_B200_PIPE_TB=$tb timeout 300 $B --workload $w 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value']/1e9, d['ms_per_step'], d['roofline']['frac'])"
done; done
HGA_B200_PIPE_TB=60 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
