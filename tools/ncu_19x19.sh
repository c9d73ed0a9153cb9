mkdir -p gpurun_out/r02
CMD="python bench.py --workload 19x19 --games 2048 --sims 100 --steps 1 --warmup 3 --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r02/plain_19.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_tc_conv|k_tree_step" -s 600 -c 4 -o gpurun_out/r02/prof_19x19 $CMD > gpurun_out/r02/ncu_f19.log 2>&1
echo ncu rc=$?; tail -3 gpurun_out/r02/ncu_f19.log; ls -la gpurun_out/r02/*.ncu-rep
