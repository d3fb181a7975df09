#!/bin/bash
TAG=${1:-r02i}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
for W in C2M C16M; do for B in 1 8; do timeout 300 python scripts/route_probe.py $W $B 2>&1 | tail -2; done; done | tee $O/route_probe.txt
timeout 600 ncu --set full --import-source on --clock-control none -k regex:shard_route -c 2 -o $O/route_C2M python scripts/route_probe.py C2M 8 > $O/ncu.log 2>&1
ncu -i $O/route_C2M.ncu-rep --page details --csv > $O/route_C2M_details.csv 2>/dev/null
grep -E "Duration|Executed Ipc Active|Theoretical Occupancy|Achieved Occupancy|Memory Throughput|DRAM Throughput|Registers Per|Block Limit|Cluster" $O/route_C2M_details.csv | cut -c1-260 | head -40
