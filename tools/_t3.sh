mkdir -p gpurun_out/r02
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_default_c.json 2> gpurun_out/r02/bench_default_c.err
echo bench rc=$?; tail -3 gpurun_out/r02/bench_default_c.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02/bench_default_c.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','mean_leaf_depth','clocks','timed_region_s')})
print(d['e2e']['value'], d['frac_step']['value'])
print({k:d['roofline'][k] for k in ('achieved','frac','launch_ms','sustained')})
PY
SEL='test_visit_counts_bit_exact_vs_reference_fixture and (N9_S50 or N6_S256) or test_recurrent_tree_equals_separate_calls and (5-33-30 or 9-64-60) or test_engine_equals_stepwise_api and (5-16-16 or 5-128-16) or test_idle_trees or test_playout_trace_bit_exact and (2-16 or 5-64 or 9-64) or test_known_answer or test_step_api or test_fp32_mode_matches_reference_golden and N6 or test_bf16_mode and N9_C128.npz or test_engine_invariants and 5-16'
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 99 python -m pytest tests -m gpu -q -x -k "$SEL" > gpurun_out/r02/sanitizer_memcheck.txt 2>&1
echo memcheck rc=$?; tail -8 gpurun_out/r02/sanitizer_memcheck.txt
