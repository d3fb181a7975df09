#!/bin/bash
# N-GPU pass: peer + routed tests (one process, GPU 0), then bench.py under torchrun as the driver launches it (both arms)
N=${1:-2}; TAG=${2:-r02k}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_routed.py tests/test_gpu_peer_exchange.py -x -q > $O/pytest_peer_routed.txt 2>&1; echo "peer+routed rc=$?"; tail -3 $O/pytest_peer_routed.txt
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
timeout 900 $L bench.py --gpus $N --steps 20 --warmup 5 $3 > $O/bench_N$N.json 2> $O/bench_N$N.err; echo "bench rc=$?"
tail -5 $O/bench_N$N.err
timeout 400 $L bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/bench_reference_N$N.json 2> $O/bench_reference_N$N.err; echo "ref rc=$?"
python - <<PY
import json
for l in open("$O/bench_N$N.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "e2e", "clocks")})[:1200])
        print(json.dumps(d["detail"])[:3500])
        print(json.dumps(d["summary"]))
for l in open("$O/bench_reference_N$N.json"):
    if l.startswith("{"):
        d = json.loads(l); print("reference:", d["value"], d["cpu_baseline"]["cores"], d["ms_per_step"])
PY
