#!/bin/bash
# First measurement of the next round: is the A65k count pass slower on a 2-GPU box?
# gpurun --gpus 2 -- 'bash scripts/gpu_trace_n2.sh r02a'
TAG=${1:-r02a}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,clocks.mem,power.draw,power.limit --format=csv > $O/smi.txt 2>&1
for D in 0 1; do
  CUDA_VISIBLE_DEVICES=$D timeout 200 python scripts/trace_hier.py A65k > $O/trace_gpu$D.txt 2>&1
  for P in 1 0; do
    CUDA_VISIBLE_DEVICES=$D HVCF_HIER_PARTS=$P timeout 300 python bench.py --workload A65k --steps 200 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline \
      > $O/bench_gpu${D}_parts$P.json 2>> $O/err.txt
  done
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 200 --warmup 5 \
  --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline > $O/bench_N2.json 2>> $O/err.txt
grep -h "pass1\|pass2" $O/trace_gpu*.txt
python - <<'PY'
import json, glob, os
for f in sorted(glob.glob(os.path.join("gpurun_out", os.environ.get("TAG", "r02a"), "bench_*.json"))):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); print(os.path.basename(f), d["n_gpus"], round(d["ms_per_step"], 5), round(d["roofline"]["kernel_ms"], 5))
PY
