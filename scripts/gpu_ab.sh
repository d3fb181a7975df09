#!/bin/bash
# A/B of two builds of libhvcf.so (HVCF_LIB): GPU suite on the default build, then per-kernel times and bench lines for both
TAG=${1:-r01zh}; ALT=${2:-hellovulkan_b200/lib/libhvcf_nopairs.so}
O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_gpu.log
for W in A65k B1M A1M; do
  for L in default alt default alt; do
    if [ $L = alt ]; then export HVCF_LIB=$PWD/$ALT; else unset HVCF_LIB; fi
    HVCF_KERNEL_TIMES=0 timeout 600 python bench.py --workload $W --steps 200 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline 2>> $O/err.txt | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$W', '$L', 'ms', round(d['ms_per_step'],5), 'count_pass_ms', round(r['kernel_ms'],5), 'frac', round(r['frac'],3), 'executed', r['executed_tests_per_launch'])
" | tee -a $O/ab.txt
  done
done
unset HVCF_LIB
timeout 600 python bench.py --workload A65k --steps 5 --warmup 3 --sharded-steps 3 2>> $O/err.txt | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('C16M one GPU', d['sharded']['ms_per_frame'])
" | tee -a $O/ab.txt
