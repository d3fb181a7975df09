#!/bin/bash
# One gpurun call: GPU tests, smoke, bench lines, ncu launch lists and full captures. Outputs under gpurun_out/$TAG.
TAG=${1:-r01c}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py > $O/bench_A65k.json 2> $O/bench_A65k.err; echo "bench A65k rc=$?" | tee -a $O/rc.txt
for W in B1M A1M A1k; do
  timeout 600 python bench.py --workload $W --steps 50 --no-sharded --no-flythrough --no-next-rows > $O/bench_$W.json 2> $O/bench_$W.err; echo "bench $W rc=$?" | tee -a $O/rc.txt
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_A65k.json 2> $O/bench_reference.err; echo "ref rc=$?" | tee -a $O/rc.txt
for W in A65k B1M; do
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$W.csv \
    python bench.py --workload $W --steps 4 --warmup 3 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline > $O/ncu_launch_$W.log 2>&1; echo "launches $W rc=$?" | tee -a $O/rc.txt
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"hier|light_transform" --launch-skip 11 -c 5 -o $O/full_$W -f \
    python bench.py --workload $W --steps 4 --warmup 3 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline > $O/ncu_full_$W.log 2>&1; echo "full $W rc=$?" | tee -a $O/rc.txt
done
cat $O/rc.txt
