#!/bin/bash
# usage: gpu_sweep.sh TAG WORKLOAD "name:ENV=V ENV=V" "name:..." ...   -- one bench.py run per variant, one summary line each
TAG=$1; W=$2; shift 2
O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
B="python bench.py --steps 20 --warmup 3 --no-summary --no-flythrough --no-next-rows --no-cpu-baseline"
for V in "$@"; do
  name=${V%%:*}; envs=${V#*:}
  env $envs timeout 300 $B --workload $W > $O/bench_${W}_$name.json 2>> $O/err.txt
  python - <<PY
import json
for l in open("$O/bench_${W}_$name.json"):
    if l.startswith("{"):
        d = json.loads(l); r = d["roofline"]
        print("$W $name", "ms", round(d["ms_per_step"], 4), {k: round(v["ms"], 4) for k, v in r["per_kernel"].items()}, "launches", d["gpu_launches_per_step"], "bin_tests", d["detail"]["stats"]["bin_tests"], "exec", d["detail"]["stats"]["executed_tests"], "pre", d["detail"]["stats"]["prefilter_tests"], "write", d["detail"]["stats"]["write_pass_tests"])
PY
done
tail -3 $O/err.txt
