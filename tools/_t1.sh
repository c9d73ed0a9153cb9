mkdir -p gpurun_out/r02
timeout 300 python -m pytest tests/test_nets_gpu.py -x -q > gpurun_out/r02/t1.txt 2>&1; echo rc=$?; tail -15 gpurun_out/r02/t1.txt
