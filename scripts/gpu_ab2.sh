#!/bin/bash
# A/B of the current library against an older build on ONE box, alternating: usage gpu_ab2.sh TAG ALT.so WORKLOAD...
TAG=$1; ALT=$2; shift 2
O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
for W in "$@"; do
  for L in new old new old new old; do
    if [ $L = old ]; then export HVCF_LIB=$PWD/$ALT; else unset HVCF_LIB; fi
    timeout 300 python bench.py --workload $W --steps 200 --warmup 5 --no-summary --no-flythrough --no-next-rows --no-cpu-baseline 2>> $O/err.txt | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$W', '$L', 'ms', round(d['ms_per_step'],5), {k: round(v['ms'],4) for k,v in r['per_kernel'].items()})
" | tee -a $O/ab.txt
  done
done
unset HVCF_LIB
