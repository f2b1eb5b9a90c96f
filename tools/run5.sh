B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$B --precision f32 > gpurun_out/plain_f32.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_hga_pipe -s 3 -c 1 -o gpurun_out/prof_pipe_f32 $B --precision f32 > gpurun_out/ncu_f32.log 2>&1
tail -2 gpurun_out/ncu_f32.log
