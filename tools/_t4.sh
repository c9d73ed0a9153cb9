mkdir -p gpurun_out/r02
timeout 600 python -m pytest tests/test_selfplay_gpu.py -x -q -k "invariant_to_sharding or engine_invariants" > gpurun_out/r02/t4.txt 2>&1; echo rc=$?; tail -5 gpurun_out/r02/t4.txt
for st in 1 2; do
python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-secondary --streams $st > gpurun_out/r02/bench_streams$st.json 2> gpurun_out/r02/bench_streams$st.err
echo bench streams=$st rc=$?; tail -2 gpurun_out/r02/bench_streams$st.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02/bench_streams$st.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','clocks')}, d['run'], d['e2e']['value'])
PY
done
python bench.py --workload 9x9 --steps 250 --warmup 5 --no-cpu-baseline --streams 2 --games 2048 > gpurun_out/r02/bench_9x9_streams2_g2048.json 2>/dev/null
python -c "
import json
d=json.loads(open('gpurun_out/r02/bench_9x9_streams2_g2048.json').read().strip().splitlines()[-1])
print('9x9 2048 games 2 streams', {k:d[k] for k in ('value','ms_per_step','mean_leaf_depth')}, d['run'])"
