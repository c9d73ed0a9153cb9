mkdir -p gpurun_out/r02
python -m muzero_go_b200.build --define MZGO_TRACE_TREE > /dev/null 2>&1; echo build rc=$?
MZGO_KEEP_SO=1 python tools/tree_trace.py > gpurun_out/r02/tree_trace_19x19.txt 2>&1; echo rc=$?; cat gpurun_out/r02/tree_trace_19x19.txt | tail -16
python -m muzero_go_b200.build --force > /dev/null 2>&1
