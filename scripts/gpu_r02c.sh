#!/bin/bash
# bench.py N=1 (new contract) + quick regression of the GPU suite pieces touched by the epoch change
TAG=${1:-r02c}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_routed.py tests/test_gpu_peer_exchange.py -x -q > $O/pytest.txt 2>&1; echo "rc=$?" >> $O/pytest.txt; tail -5 $O/pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_N1.json 2> $O/bench_N1.err; echo "bench rc=$?"
tail -3 $O/bench_N1.err
python - <<'PY'
import json
for l in open("gpurun_out/r02c/bench_N1.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "e2e", "clocks", "cpu_baseline")})[:1500])
        print(json.dumps(d["roofline"])[:1800])
        print(json.dumps(d["summary"]))
PY
