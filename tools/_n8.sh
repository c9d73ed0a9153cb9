mkdir -p gpurun_out/r02
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02/smoke.txt 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r02/smoke.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02/bench_19x19_n8.json 2> gpurun_out/r02/bench_19x19_n8.err
echo bench rc=$?; tail -3 gpurun_out/r02/bench_19x19_n8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_19x19_n8.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','n_gpus','ms_per_step','mean_leaf_depth','clocks')}, d['e2e']['value'], d['run'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step')}, s['e2e']['value'], s['from_empty_boards']['value'])
PY
