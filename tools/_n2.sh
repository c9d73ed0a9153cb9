mkdir -p gpurun_out/r02
timeout 600 python -m pytest tests/test_selfplay_gpu.py -x -q -k "invariant_to_sharding" > gpurun_out/r02/t6.txt 2>&1; echo rc=$?; tail -3 gpurun_out/r02/t6.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/r02/bench_19x19_n2.json 2> gpurun_out/r02/bench_19x19_n2.err
echo bench rc=$?; tail -3 gpurun_out/r02/bench_19x19_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_19x19_n2.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','n_gpus','ms_per_step','mean_leaf_depth','clocks')}, d['e2e']['value'], d['run'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step')}, s['e2e']['value'])
PY
