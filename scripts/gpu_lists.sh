#!/bin/bash
TAG=${1:-r01o}
O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log
for B in 0 1; do
  for W in A65k B1M A1M A1k; do
    HVCF_SEG_LISTS=$B timeout 300 python bench.py --workload $W --steps 50 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline > $O/bench_${W}_lists_$B.json 2> $O/bench_${W}_lists_$B.err; echo "$W lists=$B rc=$?" | tee -a $O/rc.txt
  done
  HVCF_SEG_LISTS=$B HVCF_KERNEL_TIMES=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-flythrough --no-next-rows --no-cpu-baseline --sharded-steps 5 > $O/sharded_N1_lists_$B.json 2> $O/sharded_N1_lists_$B.err; echo "sharded lists=$B rc=$?" | tee -a $O/rc.txt
done
