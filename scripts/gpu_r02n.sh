#!/bin/bash
TAG=${1:-r02n}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_routed.py -x -q > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.txt
B="python bench.py --steps 20 --warmup 3 --no-summary --no-flythrough --no-next-rows --no-cpu-baseline"
run() { # name, workload, env...
  local name=$1 w=$2; shift 2
  env "$@" timeout 300 $B --workload $w > $O/bench_${w}_$name.json 2>> $O/err.txt
  python - <<PY
import json
for l in open("$O/bench_${w}_$name.json"):
    if l.startswith("{"):
        d = json.loads(l); r = d["roofline"]
        print("$w $name", "ms", round(d["ms_per_step"], 4), {k: round(v["ms"], 4) for k, v in r["per_kernel"].items()}, "launches", d["gpu_launches_per_step"], "bin_tests", d["detail"]["stats"]["bin_tests"], "exec", d["detail"]["stats"]["executed_tests"], "pre", d["detail"]["stats"]["prefilter_tests"], "write", d["detail"]["stats"]["write_pass_tests"])
PY
}
run default C16M HVCF_ROUNDS=-1
run ls16 C16M HVCF_ROUNDS_LATE_SEGS=16
run ls64 C16M HVCF_ROUNDS_LATE_SEGS=64
run g2 C16M HVCF_ROUNDS_GROWTH=2
run h32g2 C16M HVCF_ROUNDS_GROWTH=2 HVCF_ROUNDS_HEAD=32
run g8 C16M HVCF_ROUNDS_GROWTH=8
run g16 C16M HVCF_ROUNDS_GROWTH=16
run B1Moff B1M HVCF_ROUNDS=0
run B1Mr3 B1M HVCF_ROUNDS=3 HVCF_ROUNDS_HEAD=32
run B1Mr2 B1M HVCF_ROUNDS=2 HVCF_ROUNDS_HEAD=64
tail -5 $O/err.txt
V=default
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 260 --csv --log-file $O/launches_C16M_$V.csv $B --steps 2 > $O/ncu_$V.log 2>&1
echo "== C16M $V"; python scripts/launch_seq.py $O/launches_C16M_$V.csv
