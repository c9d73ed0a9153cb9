mkdir -p gpurun_out/r02
rm -f gpurun_out/nets_errors.json
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/r02/pytest_gpu.txt 2>&1
echo pytest rc=$?; tail -15 gpurun_out/r02/pytest_gpu.txt
