#!/bin/bash
TAG=${1:-r02r}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_routed.py tests/test_gpu_peer_exchange.py -x -q > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.txt
python scripts/rounds_probe.py C16M 2>&1 | cut -c1-400
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
run g8 C16M HVCF_ROUNDS_GROWTH=8
run es64 C16M HVCF_ROUNDS_EARLY_SEGS=64
run es32 C16M HVCF_ROUNDS_EARLY_SEGS=32
run h32 C16M HVCF_ROUNDS_HEAD=32
run h128g8 C16M HVCF_ROUNDS_HEAD=128 HVCF_ROUNDS_GROWTH=8
run h32g2x C16M HVCF_ROUNDS_HEAD=32 HVCF_ROUNDS_GROWTH=2 HVCF_ROUNDS=4
tail -5 $O/err.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 260 --csv --log-file $O/launches_C16M.csv $B --steps 2 > $O/ncu.log 2>&1
echo "== C16M"; python scripts/launch_seq.py $O/launches_C16M.csv
