#!/bin/bash
# bench.py at the world size gpurun was given, launched the way the driver launches it
TAG=${1:-r01h}; N=${2:-8}
O=gpurun_out/$TAG; mkdir -p $O
if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"; fi
timeout 900 $L bench.py --gpus $N --steps 100 --warmup 5 > $O/bench_N$N.json 2> $O/bench_N$N.err; echo "N$N rc=$?" | tee -a $O/rc.txt
tail -n 3 $O/bench_N$N.err
