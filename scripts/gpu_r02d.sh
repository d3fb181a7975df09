#!/bin/bash
# bench.py under torchrun at N GPUs (routed frame), both arms as the driver launches them
N=${1:-2}; BENCH_ARGS=${3:-}; TAG=${2:-r02d}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 $BENCH_ARGS \
  > $O/bench_N$N.json 2> $O/bench_N$N.err; echo "bench rc=$?"
tail -5 $O/bench_N$N.err
python - <<PY
import json
for l in open("$O/bench_N$N.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "e2e", "clocks")})[:1200])
        print(json.dumps(d["detail"])[:3000])
        print(json.dumps(d["summary"]))
PY
