mkdir -p gpurun_out/r02
python -m muzero_go_b200.build --checked > gpurun_out/r02/checked_build.log 2>&1; echo build rc=$?
MZGO_KEEP_SO=1 timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02/pytest_gpu_checked_build.txt 2>&1; echo pytest rc=$?; tail -3 gpurun_out/r02/pytest_gpu_checked_build.txt; grep -c "Assertion" gpurun_out/r02/pytest_gpu_checked_build.txt
MZGO_KEEP_SO=1 python bench.py --workload 9x9 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_9x9_checked.json 2> gpurun_out/r02/bench_9x9_checked.err; echo rc9=$?; grep -c Assertion gpurun_out/r02/bench_9x9_checked.err
MZGO_KEEP_SO=1 python bench.py --workload 19x19 --games 256 --sims 400 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/r02/bench_19_checked.json 2> gpurun_out/r02/bench_19_checked.err; echo rc19=$?; grep -c Assertion gpurun_out/r02/bench_19_checked.err
python -m muzero_go_b200.build --force > /dev/null 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_optin.json 2> gpurun_out/r02/bench_optin.err; echo bench rc=$?
python bench.py --workload 9x9 --steps 250 --warmup 5 --no-cpu-baseline > gpurun_out/r02/bench_9x9_primary.json 2>/dev/null
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_optin.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','clocks')}, d['e2e']['value'], d['roofline_tree']['launch_ms'], d['roofline_tree']['frac'])
for f in ('bench_9x9_checked','bench_19_checked','bench_9x9_primary'):
    try:
        d=json.loads(open(f'gpurun_out/r02/{f}.json').read().strip().splitlines()[-1])
        print(f, d['value'], {k:d['roofline_tree'][k] for k in ('launch_ms','frac','mean_leaf_depth','sims_per_launch')})
    except Exception as e:
        print(f, 'no line', e)
PY
