#!/bin/bash
# round 2, first call: full-size oracle parity + the A65k count pass bimodality on ONE box (fresh processes)
TAG=${1:-r02a}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,clocks.mem,power.draw,power.limit --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
timeout 900 python -m pytest tests/test_gpu_routed.py -x -q --durations=5 > $O/pytest_routed.txt 2>&1
echo "routed rc=$?" >> $O/pytest_routed.txt
tail -25 $O/pytest_routed.txt
timeout 1500 python -m pytest tests/test_gpu_full_size.py -x -q --durations=10 > $O/pytest_full_size.txt 2>&1
echo "full-size rc=$?" >> $O/pytest_full_size.txt
tail -15 $O/pytest_full_size.txt
for I in 1 2 3 4 5 6; do
  timeout 200 python scripts/trace_hier.py A65k > $O/trace_$I.txt 2>&1
  timeout 300 python bench.py --workload A65k --steps 200 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline \
      > $O/bench_$I.json 2>> $O/err.txt
done
grep -h "pass1" $O/trace_*.txt
python - <<'PY'
import json, glob, os
for f in sorted(glob.glob("gpurun_out/*/bench_*.json")):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); print(os.path.basename(f), round(d["ms_per_step"], 5), round(d["roofline"]["kernel_ms"], 5), d["roofline"]["executed_tests_per_launch"], d["clocks"])
PY
