#!/bin/bash
# Round profile on one B200 (run under gpurun from the repo root): tests, both bench workloads, the CPU arm,
# ncu launch lists and --set full captures of the two dominant kernels.  Outputs land in gpurun_out/.
set -x
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
python bench.py --steps 8 --warmup 3 > $O/bench_9x9.json 2> $O/bench_9x9.err
python bench.py --workload 19x19 --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_19x19.json 2> $O/bench_19x19.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_9x9_reference.json 2> $O/bench_9x9_reference.err
# launch lists (shares of the step; per-launch times under ncu are cold-cache and serialised)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/plain_9x9.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 420 --csv --log-file $O/launches_9x9.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_l9.log 2>&1
python bench.py --workload 19x19 --sims 100 --steps 1 --warmup 3 --no-cpu-baseline > $O/plain_19x19.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 420 --csv --log-file $O/launches_19x19_s100.csv \
    python bench.py --workload 19x19 --sims 100 --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_l19.log 2>&1
# top kernels, --set full (same commands as the plain runs just above)
ncu --set full --clock-control none --import-source on -k regex:"k_tc_conv|k_tree_step" -s 300 -c 4 -o $O/prof_9x9 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_f9.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_tc_conv|k_tree_step" -s 600 -c 4 -o $O/prof_19x19 \
    python bench.py --workload 19x19 --sims 100 --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_f19.log 2>&1
ls -la $O
