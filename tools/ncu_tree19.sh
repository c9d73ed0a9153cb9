# ncu --set full source-level capture of the 19x19 tree step deep inside a search (PUCT at the root, level 2 below):
# launch 600 of the second move.  Run under gpurun from the repo root; plain run first.
mkdir -p gpurun_out/r02
CMD="python bench.py --workload 19x19 --games 1024 --sims 800 --steps 1 --warmup 3 --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r02/plain_tree19.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_tree_step" -s 1400 -c 2 -o gpurun_out/r02/prof_tree19 $CMD > gpurun_out/r02/ncu_tree19.log 2>&1
echo ncu rc=$?; tail -3 gpurun_out/r02/ncu_tree19.log; tail -2 gpurun_out/r02/plain_tree19.log | cut -c1-600; ls -la gpurun_out/r02/*.ncu-rep
