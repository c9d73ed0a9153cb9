#!/bin/bash
# What the driver runs at round end, on one B200: GPU tests, smoke, the default bench line, the 9x9 workload as primary.
O=gpurun_out/r02
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu_final.txt 2>&1; echo pytest rc=$?; tail -3 $O/pytest_gpu_final.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo smoke rc=$?; tail -1 $O/smoke.txt
python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_19x19_default_steps20.json 2> $O/bench_19x19_default_steps20.err; echo bench rc=$?
python bench.py --workload 9x9 --steps 250 --warmup 5 --no-cpu-baseline > $O/bench_9x9_primary.json 2> $O/bench_9x9_primary.err; echo bench9 rc=$?
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_19x19_default_steps20.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','moves_per_sec','clocks','gpu_launches')})
print(d['e2e'], d['frac_step']['value'], d['cpu_baseline']['value'])
print({k:d['roofline'][k] for k in ('achieved','frac','launch_ms')}, d['roofline']['sustained']['frac'], d['roofline_tree']['launch_ms'], d['roofline_tree']['frac'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step')}, s['from_empty_boards']['value'])
d=json.loads(open('gpurun_out/r02/bench_9x9_primary.json').read().strip().splitlines()[-1])
print('9x9 primary', {k:d[k] for k in ('value','ms_per_step','mean_leaf_depth')}, d['e2e']['value'], d['roofline']['frac'], d['roofline_tree']['frac'], d['roofline_rules']['frac'])
PY
