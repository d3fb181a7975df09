#!/bin/bash
# the e2e leg with two frames in flight: the headline workload alone (no summary legs), then the driver's full default line
TAG=${1:-r04b}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 400 python bench.py --no-summary --no-flythrough --no-next-rows --no-cpu-baseline > $O/bench_head.json 2> $O/bench_head.err; echo "head rc=$?" | tee -a $O/rc.txt
tail -5 $O/bench_head.err
python - <<PY
import json
for l in open("$O/bench_head.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "e2e")})[:2500])
PY
cat $O/rc.txt
