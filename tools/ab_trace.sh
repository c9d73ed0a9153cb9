#!/bin/bash
# tools/ab_trace.sh "<target>=<file>,..." ...   : builds each variant on the box and prints the 19x19 role trace summary
mkdir -p gpurun_out/ab/keep
for spec in "$@"; do
  for kv in ${spec//,/ }; do t=${kv%%=*}; f=${kv#*=}; [ -f gpurun_out/ab/keep/$(basename $t) ] || cp $t gpurun_out/ab/keep/$(basename $t); cp $f $t; done
  python -m muzero_go_b200.build > gpurun_out/ab/build.log 2>&1 || tail -5 gpurun_out/ab/build.log
  echo "== $spec"
  python tools/tc_trace.py --n 19 --leaves 1024 2>&1 | grep "per launch\|group 0 start\|group 6 issued\|leave"
  for kv in ${spec//,/ }; do t=${kv%%=*}; cp gpurun_out/ab/keep/$(basename $t) $t; done
done
python -m muzero_go_b200.build > /dev/null 2>&1
