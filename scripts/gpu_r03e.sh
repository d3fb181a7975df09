#!/bin/bash
# round 2, final single-GPU evidence pass: GPU tests, smoke, the driver's bench line (C16M) + reference arm, ncu launch list and
# full capture of the C16M frame's kernels, rounds on/off on the 1M-light workloads.  Outputs under gpurun_out/$TAG.
TAG=${1:-r03e}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -4 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py > $O/bench_N1.json 2> $O/bench_N1.err; echo "bench rc=$?" | tee -a $O/rc.txt
tail -3 $O/bench_N1.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?" | tee -a $O/rc.txt
ARGS="--steps 2 --warmup 3 --no-summary --no-flythrough --no-next-rows --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/launches_C16M.csv \
  python bench.py $ARGS > $O/ncu_launch.log 2>&1; echo "launches rc=$?" | tee -a $O/rc.txt
python scripts/launch_seq.py $O/launches_C16M.csv | tee $O/launch_seq_C16M.txt
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"hier|light_transform" --launch-skip 34 -c 17 -o $O/full_C16M -f \
  python bench.py $ARGS > $O/ncu_full.log 2>&1; echo "full rc=$?" | tee -a $O/rc.txt
ncu -i $O/full_C16M.ncu-rep --page raw --csv > $O/full_C16M_raw.csv 2>/dev/null
python scripts/ncu_summary.py $O/full_C16M_raw.csv > $O/full_C16M_summary.txt 2>&1
cut -c1-260 $O/full_C16M_summary.txt
rm -f $O/full_C16M.ncu-rep
python - <<PY
import json
for l in open("$O/bench_N1.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "e2e", "clocks", "cpu_baseline")})[:1500])
        print(json.dumps(d["roofline"])[:2500])
        print(json.dumps(d["summary"]))
PY
cat $O/rc.txt
