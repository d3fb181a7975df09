#!/bin/bash
TAG=${1:-r03k}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 300 python scripts/route_probe.py C16M 16 2>&1 | tail -3 | tee $O/route_probe.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:shard_route --launch-skip 5 -c 1 -o $O/route -f python scripts/route_probe.py C16M 16 > $O/ncu.log 2>&1
ncu -i $O/route.ncu-rep --page raw --csv > $O/route_raw.csv 2>/dev/null
python scripts/ncu_summary.py $O/route_raw.csv | cut -c1-900
ncu -i $O/route.ncu-rep --page details --csv 2>/dev/null | grep -E "Registers Per|Theoretical Occupancy|Achieved Occupancy|Block Limit|Cluster|Duration|Executed Ipc|No Eligible|Stall" | cut -c1-300 | head -40
ncu -i $O/route.ncu-rep --page source --csv --print-source sass 2>/dev/null > $O/route_source.csv; wc -l $O/route_source.csv
rm -f $O/route.ncu-rep
