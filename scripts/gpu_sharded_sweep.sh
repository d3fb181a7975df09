#!/bin/bash
# sweep of the pipelined light exchange at the world size given: chunks x NCCL stream priority
TAG=${1:-r01g}; N=${2:-2}
O=gpurun_out/$TAG; mkdir -p $O
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for C in 4 8 16 32; do
  timeout 300 $L bench.py --gpus $N --steps 5 --warmup 3 --no-flythrough --no-next-rows --no-cpu-baseline --sharded-steps 5 --sharded-chunks $C > $O/N${N}_c$C.json 2> $O/N${N}_c$C.err; echo "c$C rc=$?" >> $O/rc.txt
done
HVCF_NCCL_LOW_PRIORITY=1 timeout 300 $L bench.py --gpus $N --steps 5 --warmup 3 --no-flythrough --no-next-rows --no-cpu-baseline --sharded-steps 5 --sharded-chunks 8 > $O/N${N}_c8_lowprio.json 2> $O/N${N}_c8_lowprio.err; echo "lowprio rc=$?" >> $O/rc.txt
cat $O/rc.txt
