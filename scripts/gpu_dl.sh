#!/bin/bash
# hvcf_download_grid with grid_download_kernel (pinned destinations) against the two-copy path: GPU suite, then e2e lines
TAG=${1:-r01ze}
O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_gpu.log
for W in A65k B1M; do
  for CFG in "0 64" "1 32" "1 64" "1 148" "1 296" "0 64" "1 64"; do
    set -- $CFG
    HVCF_DOWNLOAD_KERNEL=$1 HVCF_DOWNLOAD_BLOCKS=$2 timeout 600 python bench.py --workload $W --steps 200 --no-sharded --no-flythrough --no-next-rows --no-cpu-baseline > $O/tmp.json 2>> $O/bench_$W.err
    echo "bench $W kernel=$1 blocks=$2 rc=$?" >> $O/rc.txt
    python - "$W" "$1" "$2" $O/tmp.json <<'PY' | tee -a $O/e2e_download.txt
import json,sys
for line in open(sys.argv[4]):
    if line.startswith("{"):
        d=json.loads(line); e=d["e2e"]
        print(sys.argv[1], "download_kernel", sys.argv[2], "blocks", sys.argv[3], "ms_per_step", round(d["ms_per_step"],5), "e2e_ms", round(e["ms_per_step"],5),
              "one_at_a_time", round(e.get("one_frame_at_a_time",{}).get("ms_per_step",0),5), "d2h", e["d2h_bytes_per_step"])
PY
  done
done
