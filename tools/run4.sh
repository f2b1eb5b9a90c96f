python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variants or f32 or chunked or k1_shape or 2khz or segments" 2>&1 | tail -3
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline"
for cfg in "f32 4 40 9" "f32 4 20 9" "f32 2 40 9" "f64 4 40 9" "f64 4 20 9" "f64 2 40 9"; do set -- $cfg; echo "prec=$1 S=$2 TB=$3 OCC=$4"; HGA_B200_PIPE_S=$2 HGA_B200_PIPE_TB=$3 HGA_B200_PIPE_OCC=$4 $B --precision $1 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['warmup_samples'])"; done
