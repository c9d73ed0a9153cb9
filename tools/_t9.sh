mkdir -p gpurun_out/r02
timeout 1200 python -m pytest tests/test_train_gpu.py tests/test_selfplay_gpu.py tests/test_arena.py -x -q -m gpu > gpurun_out/r02/t9.txt 2>&1; echo rc=$?; tail -4 gpurun_out/r02/t9.txt
