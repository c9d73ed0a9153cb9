mkdir -p gpurun_out/r02
python -m muzero_go_b200.build --checked > gpurun_out/r02/checked_build.log 2>&1; echo build rc=$?
MZGO_KEEP_SO=1 timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02/pytest_gpu_checked_build.txt 2>&1; echo pytest rc=$?
tail -4 gpurun_out/r02/pytest_gpu_checked_build.txt
grep -c "Assertion" gpurun_out/r02/pytest_gpu_checked_build.txt
python -m muzero_go_b200.build --force > /dev/null 2>&1
