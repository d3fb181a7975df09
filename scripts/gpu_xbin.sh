#!/bin/bash
# fused transform + binning (hier_xbin_kernel) against the two kernels: GPU suite, then bench lines with the fusion on / off
TAG=${1:-r01zc}
O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_gpu.log
for W in A65k A1k; do
  for MX in 0 262144 0 262144; do
    HVCF_XBIN_MAX=$MX timeout 600 python bench.py --workload $W --steps 200 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline >> $O/bench_${W}_xbin$MX.json 2>> $O/bench_$W.err
    echo "bench $W xbin_max=$MX rc=$?" | tee -a $O/rc.txt
  done
done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob(os.path.join("gpurun_out", os.environ.get("TAG","r01zc"), "bench_*_xbin*.json"))):
    for line in open(f):
        line=line.strip()
        if not line.startswith("{"): continue
        d=json.loads(line)
        print(os.path.basename(f), d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("gpu_launches"), d.get("kernel_ms"))
PY
