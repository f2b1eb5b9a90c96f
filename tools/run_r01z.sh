B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-stream"
for w in k2 k4; do
echo "== $w"; timeout 300 $B --workload $w 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['other_precisions'])"
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
