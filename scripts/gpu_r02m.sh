#!/bin/bash
TAG=${1:-r02m}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
B="python bench.py --steps 2 --warmup 3 --no-summary --no-flythrough --no-next-rows --no-cpu-baseline"
for V in default r3; do
  E=""; [ $V = r3 ] && E="HVCF_ROUNDS=3"
  env $E timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 260 --csv --log-file $O/launches_C16M_$V.csv $B > $O/ncu_$V.log 2>&1
  echo "== C16M $V"; python scripts/launch_seq.py $O/launches_C16M_$V.csv
done
