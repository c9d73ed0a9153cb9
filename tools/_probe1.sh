mkdir -p gpurun_out/r02
{
echo "== gym_go probe $(date -u)"
python -c "import gym_go; print('gym_go importable', gym_go.__file__)" 2>&1 | tail -1
python -c "import gym; print('gym importable', gym.__file__)" 2>&1 | tail -1
python -c "import gymnasium; print('gymnasium importable', gymnasium.__file__)" 2>&1 | tail -1
timeout 60 python -m pip download --no-deps -d /tmp/dl gym 2>&1 | tail -2
timeout 60 python -m pip install --no-index --find-links /opt/wheelhouse gym 2>&1 | tail -2
ls /opt/wheelhouse 2>/dev/null | grep -i -E "gym|go" 
timeout 60 git clone --depth 1 https://github.com/huangeddie/GymGo /tmp/GymGo 2>&1 | tail -2
find / -xdev \( -iname "gym_go*" -o -iname "GymGo*" -o -iname "gogame*" \) -not -path "/proc/*" 2>/dev/null | head
ls baseline/_ref 2>&1 | head -3
echo "== end probe"
} > gpurun_out/r02/gymgo_probe.txt 2>&1
nvidia-smi --query-gpu=name,power.limit,clocks.max.sm --format=csv >> gpurun_out/r02/gymgo_probe.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/pytest_gpu_1.txt 2>&1
echo pytest rc=$?; tail -5 gpurun_out/r02/pytest_gpu_1.txt
python bench.py --workload 19x19 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02/bench_19x19_r01head.json 2> gpurun_out/r02/bench_19x19_r01head.err
echo bench rc=$?
tail -c 600 gpurun_out/r02/bench_19x19_r01head.json
