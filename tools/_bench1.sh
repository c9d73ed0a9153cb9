mkdir -p gpurun_out/r02
python bench.py --steps 6 --warmup 3 > gpurun_out/r02/bench_default_a.json 2> gpurun_out/r02/bench_default_a.err
echo bench rc=$?; tail -3 gpurun_out/r02/bench_default_a.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_default_a.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','moves_per_sec','clocks','timed_region_s')})
print(d['e2e'], d['frac_step']['value'])
print({k:d['roofline'][k] for k in ('achieved','frac','frac_of_sustained','launch_ms','traffic')})
print({k:d['roofline_tree'][k] for k in ('achieved','frac','launch_ms','mean_leaf_depth')})
print({k:d['roofline_rules'][k] for k in ('achieved','frac','launch_ms','cpu_baseline')})
print(d['cpu_baseline'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step','mean_leaf_depth','frac_step')}, s['e2e']['value'], s['roofline']['frac'])
PY
