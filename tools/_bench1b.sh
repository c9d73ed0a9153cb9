mkdir -p gpurun_out/r02
gpurun_out/mma2_probe > gpurun_out/r02/mma2_probe.txt 2>&1; echo probe rc=$?; cat gpurun_out/r02/mma2_probe.txt
bash tools/_bench1.sh
