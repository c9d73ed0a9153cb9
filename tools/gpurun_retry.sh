#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit 3, nothing charged).  usage: tools/gpurun_retry.sh <timeout> '<command>' [gpus]
T=$1; CMD=$2; N=${3:-1}
for i in $(seq 1 12); do
  if [ "$N" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $T -- "$CMD"; else /usr/local/graft/bin/gpurun --gpus $N --timeout $T -- "$CMD"; fi
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[retry] busy, attempt $i; sleeping 90 s"; sleep 90
done
exit 3
