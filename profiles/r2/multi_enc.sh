#!/bin/bash
# N concurrent Thorenc_gpu processes on one GPU (optionally under CUDA MPS): aggregate frames/s.  usage: multi_enc.sh <N> <config> <w> <h> <frames> <mps 0|1>
N=$1; CFG=$2; W=$3; H=$4; F=$5; MPS=$6
CMD=$(python profiles/r2/enc_cmd.py $CFG $W $H $F 32)
if [ "$MPS" = 1 ]; then
  export CUDA_MPS_PIPE_DIRECTORY=/tmp/mps_pipe CUDA_MPS_LOG_DIRECTORY=/tmp/mps_log
  mkdir -p $CUDA_MPS_PIPE_DIRECTORY $CUDA_MPS_LOG_DIRECTORY
  nvidia-cuda-mps-control -d || { echo "mps daemon failed"; exit 3; }
  sleep 1
fi
T0=$(date +%s.%N)
for i in $(seq 1 $N); do
  ( THORCUDA_HOOK_TIMING=1 ${CMD//\/dev\/shm\/p/\/dev\/shm\/p$i} > /dev/shm/multi_$i.log 2>&1 ) &
done
wait
T1=$(date +%s.%N)
if [ "$MPS" = 1 ]; then echo quit | nvidia-cuda-mps-control; fi
python - <<PY
import re,glob
n=$N; f=$F; dt=$T1-$T0
st=[float(m.group(1)) for p in glob.glob('/dev/shm/multi_*.log') for m in re.finditer(r'= ([0-9.]+) frames/s', open(p).read())]
print({"procs": n, "mps": $MPS, "config": "$CFG", "size": "${W}x${H}", "frames_each": f, "wall_s": round(dt,2), "aggregate_fps_wall": round(n*f/dt,2), "steady_fps_each": st, "steady_fps_sum": round(sum(st),2)})
PY
