mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_selfplay_gpu.py tests/test_nets_gpu.py -x -q > gpurun_out/r02/t5.txt 2>&1; echo rc=$?; tail -4 gpurun_out/r02/t5.txt
for st in 1 2; do
python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-secondary --streams $st > gpurun_out/r02/bench_co_streams$st.json 2> gpurun_out/r02/bench_co_streams$st.err
echo bench streams=$st rc=$?; tail -2 gpurun_out/r02/bench_co_streams$st.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02/bench_co_streams$st.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','clocks')}, d['run']['waves'], d['e2e']['value'], d['roofline']['launch_ms'], d['roofline']['sustained']['launch_ms'], d['roofline_tree']['launch_ms'])
PY
done
for g in 1024 2048; do
python bench.py --workload 9x9 --steps 250 --warmup 5 --no-cpu-baseline --streams 2 --games $g > gpurun_out/r02/bench_9x9_co_streams2_g$g.json 2>/dev/null
python -c "
import json
d=json.loads(open('gpurun_out/r02/bench_9x9_co_streams2_g$g.json').read().strip().splitlines()[-1])
print('9x9 $g games 2 streams', {k:d[k] for k in ('value','ms_per_step','mean_leaf_depth')}, d['run']['waves'], d['roofline']['launch_ms'])"
done
python bench.py --workload 9x9 --steps 250 --warmup 5 --no-cpu-baseline --streams 1 > gpurun_out/r02/bench_9x9_co_streams1.json 2>/dev/null
python -c "
import json
d=json.loads(open('gpurun_out/r02/bench_9x9_co_streams1.json').read().strip().splitlines()[-1])
print('9x9 1024 games 1 stream', {k:d[k] for k in ('value','ms_per_step','mean_leaf_depth')}, d['run']['waves'], d['roofline']['launch_ms'])"
