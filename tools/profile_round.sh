#!/bin/bash
# Round profile on one B200 (run under gpurun from the repo root): ncu launch lists (shares of a step) and --set full
# captures of the dominant kernels.  Every ncu command runs only after the same command exited 0 without ncu.
# Outputs land in gpurun_out/r02/; tools/summarise_ncu.py turns them into the tracked files under profiles/r02/.
O=gpurun_out/r02
mkdir -p $O
C19="python bench.py --workload 19x19 --games 2048 --sims 100 --steps 1 --warmup 3 --no-cpu-baseline --no-secondary"
C9="python bench.py --workload 9x9 --steps 2 --warmup 3 --no-cpu-baseline"
# launch lists (per-launch times under ncu are cold-cache and serialised: compare SHARES)
$C19 > $O/plain_19.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 460 --csv --log-file $O/launches_19x19_s100.csv $C19 > $O/ncu_l19.log 2>&1
echo "launches 19 rc=$?"
$C9 > $O/plain_9.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 420 --csv --log-file $O/launches_9x9.csv $C9 > $O/ncu_l9.log 2>&1
echo "launches 9 rc=$?"
# --set full: 9x9 fused kernel, rules kernels (19x19 conv + tree step: tools/ncu_19x19.sh, same command as C19)
ncu --set full --clock-control none --import-source on -k regex:"k_tc_conv" -s 300 -c 2 -o $O/prof_9x9 $C9 > $O/ncu_f9.log 2>&1
echo "full 9 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_rules_playout" -c 12 -o $O/prof_rules19 $C19 > $O/ncu_fr.log 2>&1
echo "full rules rc=$?"
ls -la $O | grep -E "csv|ncu-rep"
