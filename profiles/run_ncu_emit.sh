CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1"
$CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:kvq_emit_kernel -s 3 -c 1 -o gpurun_out/prof_emit_v5b -f $CMD > gpurun_out/ncu_emit_v5b.log 2>&1
tail -2 gpurun_out/ncu_emit_v5b.log
