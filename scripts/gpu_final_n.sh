#!/bin/bash
# final N-GPU bench lines, both arms, launched as the driver launches them
N=${1:-2}; TAG=${2:-r03z}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
nvidia-smi topo -m > $O/topo_N$N.txt 2>&1
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
timeout 400 $L bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/bench_reference_N$N.json 2> $O/bench_reference_N$N.err; echo "ref rc=$?"
timeout 900 $L bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_N$N.json 2> $O/bench_N$N.err; echo "bench rc=$?"
tail -3 $O/bench_N$N.err
python - <<PY
import json
for l in open("$O/bench_N$N.json"):
    if l.startswith("{"):
        d = json.loads(l); t = d["detail"]
        print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "clocks")}), "e2e", d["e2e"]["ms_per_step"], "res", d["e2e"]["lights_resident"]["ms_per_step"])
        print("cuts", t["row_cuts"], "own", t["own_kernels_ms_per_rank"], "inbox", t["inbox_records_per_rank"], "equal", t["sharded_bytes_equal"])
        for k in t["kernel_ms_per_rank"]: print("    ", k)
        print(json.dumps(d["summary"]))
for l in open("$O/bench_reference_N$N.json"):
    if l.startswith("{"):
        d = json.loads(l); print("reference:", d["value"], d["cpu_baseline"]["cores"], d["ms_per_step"])
PY
