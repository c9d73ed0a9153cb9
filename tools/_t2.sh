mkdir -p gpurun_out/r02
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02/pytest_gpu.txt 2>&1; echo pytest rc=$?; tail -5 gpurun_out/r02/pytest_gpu.txt
python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_default_b.json 2> gpurun_out/r02/bench_default_b.err
echo bench rc=$?; tail -3 gpurun_out/r02/bench_default_b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_default_b.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','moves_per_sec','clocks','timed_region_s')})
print(d['e2e'], d['frac_step']['value'])
print({k:d['roofline'][k] for k in ('achieved','frac','frac_of_sustained','launch_ms','traffic')})
print({k:d['roofline_tree'][k] for k in ('achieved','frac','launch_ms','mean_leaf_depth')})
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step','mean_leaf_depth','frac_step')}, s['e2e']['value'], s['roofline']['frac'], s['roofline']['launch_ms'])
PY
