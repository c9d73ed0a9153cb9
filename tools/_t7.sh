mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_train_gpu.py tests/test_arena.py -x -q -m gpu > gpurun_out/r02/t7.txt 2>&1; echo rc=$?; tail -15 gpurun_out/r02/t7.txt
python bench.py --train --steps 20 --warmup 3 > gpurun_out/r02/bench_train_native.json 2> gpurun_out/r02/bench_train_native.err; echo train rc=$?; tail -2 gpurun_out/r02/bench_train_native.err; cat gpurun_out/r02/bench_train_native.json | cut -c1-900
