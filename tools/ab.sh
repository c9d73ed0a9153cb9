#!/bin/bash
# Same-box A/B of source variants (box-to-box noise of the bench line is ~1 %):
#   tools/ab.sh "<target>=<variant file>[,<target>=<variant file>...]" "<...for B>" [bench args]
# builds each variant on the GPU box, runs bench.py alternately (A B A B) and prints value / ms per step / tree step.
VA=$1; VB=$2; shift 2
ARGS=${@:---no-cpu-baseline --no-secondary --steps 10}
mkdir -p gpurun_out/ab/keep
apply() { for kv in ${1//,/ }; do t=${kv%%=*}; f=${kv#*=}; [ -f gpurun_out/ab/keep/$(basename $t) ] || cp $t gpurun_out/ab/keep/$(basename $t); cp $f $t; done; }
restore() { for kv in ${1//,/ }; do t=${kv%%=*}; cp gpurun_out/ab/keep/$(basename $t) $t; done; }
for round in 1 2; do
  for v in A B; do
    spec=$VA; [ $v = B ] && spec=$VB
    apply $spec
    python -m muzero_go_b200.build > gpurun_out/ab/build_$v.log 2>&1 || { echo "build $v failed"; tail -5 gpurun_out/ab/build_$v.log; }
    python bench.py $ARGS > gpurun_out/ab/bench_${v}${round}.json 2> gpurun_out/ab/bench_${v}${round}.err
    python - <<PY
import json
d = json.loads(open("gpurun_out/ab/bench_${v}${round}.json").read().strip().splitlines()[-1])
rt = d.get("roofline_tree") or {}
print("$v$round", round(d["value"] / 1e6, 3), "M  ms/step", round(d["ms_per_step"], 2), " tree alone us", round(1e3 * rt.get("launch_ms", 0), 2),
      " net alone us", round(1e3 * d["roofline"]["launch_ms"], 2), " sustained us", round(1e3 * d["roofline"]["sustained"]["launch_ms"], 2))
PY
    restore $spec
  done
done
python -m muzero_go_b200.build > /dev/null 2>&1
