#!/bin/bash
# one scaling point on an N-GPU box: the default bench (as the driver launches it) and the BASELINE config 4 render-only line
# usage (under gpurun --gpus N): bash tools/scale_run.sh N TAG
N=$1; TAG=$2
if [ "$N" = "1" ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571"; fi
$RUN bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err
echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench_${N}gpu.err
$RUN bench.py --gpus $N --workload C4 --render-only --steps 16 --no-cpu > gpurun_out/${TAG}_c4_${N}gpu.json 2> gpurun_out/${TAG}_c4_${N}gpu.err
echo "c4 rc=$?"; tail -c 200 gpurun_out/${TAG}_c4_${N}gpu.err
head -c 400 gpurun_out/${TAG}_c4_${N}gpu.json
