set -x
nvidia-smi --query-gpu=index,name --format=csv
timeout 600 python bench.py --workload k4 --steps 10 --warmup 3 --no-e2e --no-stream --no-cpu-baseline > gpurun_out/bench_k4.json 2> gpurun_out/bench_k4.err; tail -1 gpurun_out/bench_k4.json | cut -c1-900
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -1 gpurun_out/bench_n2.json | cut -c1-1200; tail -3 gpurun_out/bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-400
