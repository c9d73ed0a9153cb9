mkdir -p gpurun_out/r02
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02/pytest_gpu_final.txt 2>&1; echo pytest rc=$?; tail -3 gpurun_out/r02/pytest_gpu_final.txt
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02/bench_19x19_driver_args.json 2> gpurun_out/r02/bench_19x19_driver_args.err; echo bench rc=$?; tail -2 gpurun_out/r02/bench_19x19_driver_args.err
python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > gpurun_out/r02/bench_19x19_reference_cpu.json 2> gpurun_out/r02/bench_19x19_reference_cpu.err; echo ref rc=$?
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_19x19_driver_args.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','moves_per_sec','clocks','timed_region_s','gpu_launches')})
print(d['e2e'], d['frac_step']['value'], d['cpu_baseline']['value'])
print({k:d['roofline'][k] for k in ('achieved','frac','launch_ms','traffic')}, d['roofline']['sustained'])
print({k:d['roofline_tree'][k] for k in ('achieved','frac','launch_ms','mean_leaf_depth','traffic')})
print({k:d['roofline_rules'][k] for k in ('achieved','frac','launch_ms','traffic')}, d['roofline_rules']['cpu_baseline']['value'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step','mean_leaf_depth','frac_step')}, s['e2e']['value'], s['roofline']['frac'], s['openings'])
r=json.loads(open('gpurun_out/r02/bench_19x19_reference_cpu.json').read().strip().splitlines()[-1])
print('reference', r['value'], r['moves_per_sec'], r['config']==d['config'])
PY
