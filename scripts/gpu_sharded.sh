#!/bin/bash
# sharded leg (configs[3]) at the world size gpurun was given: plain + HVCF_KERNEL_TIMES=1
TAG=${1:-r01d}; N=${2:-1}
O=gpurun_out/$TAG; mkdir -p $O
ARGS="--steps 20 --warmup 3 --no-flythrough --no-next-rows --no-cpu-baseline --sharded-steps 5 $EXTRA_ARGS"
if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"; fi
timeout 600 $L bench.py --gpus $N $ARGS > $O/bench_N$N.json 2> $O/bench_N$N.err; echo "N$N rc=$?" | tee -a $O/rc.txt
HVCF_KERNEL_TIMES=1 timeout 600 $L bench.py --gpus $N $ARGS > $O/bench_N${N}_ktimes.json 2> $O/bench_N${N}_ktimes.err; echo "N$N ktimes rc=$?" | tee -a $O/rc.txt
tail -c 2500 $O/bench_N$N.json
