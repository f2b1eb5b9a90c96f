set -x
nvidia-smi --query-gpu=index,name --format=csv
for n in 2 4 8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_f_n$n.json 2> gpurun_out/bench_f_n$n.err; tail -1 gpurun_out/bench_f_n$n.json | cut -c1-400
done
