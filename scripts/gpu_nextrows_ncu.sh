#!/bin/bash
# the rows next to the path: bench figures + one ncu --set full capture of cluster_shade_kernel and frustum_cull_kernel
TAG=${1:-r03q}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
B="python bench.py --workload A65k --steps 3 --warmup 3 --no-summary --no-flythrough --no-cpu-baseline"
timeout 600 $B > $O/bench_next_rows.json 2> $O/err.txt
python - <<PY
import json
for l in open("$O/bench_next_rows.json"):
    if l.startswith("{"):
        d = json.loads(l)["next_rows"]
        print(json.dumps(d["consumer"])[:700]); print(json.dumps(d["frustum"])[:1200])
PY
timeout 900 ncu --set full --clock-control none -k regex:"cluster_shade|frustum_cull" --launch-skip 3 -c 12 -o $O/next_rows -f $B > $O/ncu.log 2>&1
ncu -i $O/next_rows.ncu-rep --page raw --csv > $O/next_rows_raw.csv 2>/dev/null
python scripts/ncu_summary.py $O/next_rows_raw.csv | cut -c1-500
rm -f $O/next_rows.ncu-rep
