mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_tree_gpu.py tests/test_selfplay_gpu.py -x -q > gpurun_out/r02/t10.txt 2>&1; echo rc=$?; tail -3 gpurun_out/r02/t10.txt
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_prewait.json 2> gpurun_out/r02/bench_prewait.err; echo bench rc=$?
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_prewait.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','clocks')}, d['e2e']['value'], d['roofline_tree']['launch_ms'], d['roofline']['sustained']['launch_ms'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step')}, s['from_empty_boards']['value'])
PY
bash tools/_trace.sh
