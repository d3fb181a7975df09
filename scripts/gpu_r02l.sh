#!/bin/bash
# rounds (hier_alive_kernel): parity + timing of C16M / B1M / A1M with rounds off / on
TAG=${1:-r02l}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.txt
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
run off C16M HVCF_ROUNDS=0
run default C16M HVCF_ROUNDS=-1
run h32g4 C16M HVCF_ROUNDS_HEAD=32
run h128g4 C16M HVCF_ROUNDS_HEAD=128
run h64g8 C16M HVCF_ROUNDS_GROWTH=8
run h64g2 C16M HVCF_ROUNDS_GROWTH=2
run r2 C16M HVCF_ROUNDS=2
run r3 C16M HVCF_ROUNDS=3
for W in B1M A1M; do
run off $W HVCF_ROUNDS=0
run r2h64 $W HVCF_ROUNDS=2 HVCF_ROUNDS_HEAD=64
run r3h32 $W HVCF_ROUNDS=3 HVCF_ROUNDS_HEAD=32
run r4h16 $W HVCF_ROUNDS=4 HVCF_ROUNDS_HEAD=16
run r8h8g2 $W HVCF_ROUNDS=8 HVCF_ROUNDS_HEAD=8 HVCF_ROUNDS_GROWTH=2
done
tail -5 $O/err.txt
