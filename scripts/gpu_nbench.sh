#!/bin/bash
# usage: gpu_nbench.sh N TAG "name:bench args" ...  -- bench.py under torchrun at N GPUs, one run per variant (routed frame only)
N=$1; TAG=$2; shift 2
O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
for V in "$@"; do
  name=${V%%:*}; args=${V#*:}
  timeout 600 $L bench.py --gpus $N --steps 20 --warmup 5 --no-compare --no-flythrough $args > $O/bench_N${N}_$name.json 2> $O/bench_N${N}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
for l in open("$O/bench_N${N}_$name.json"):
    if l.startswith("{"):
        d = json.loads(l); t = d["detail"]
        print("N$N $name ms", round(d["ms_per_step"], 4), "equal", t["sharded_bytes_equal"], "cuts", t["row_cuts"], "own", t["own_kernels_ms_per_rank"], "route", t["route_kernel_ms_per_rank"], "inboxM", [round(x / 1e6, 2) for x in t["inbox_records_per_rank"]])
        for k in t["kernel_ms_per_rank"]: print("    ", k)
PY
done
