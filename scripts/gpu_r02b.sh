#!/bin/bash
# routed frame: correctness on one GPU (several shard contexts in one process)
TAG=${1:-r02b}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 1200 python -m pytest tests/test_gpu_routed.py -x -q --durations=5 > $O/pytest_routed.txt 2>&1
echo "routed rc=$?" >> $O/pytest_routed.txt
tail -40 $O/pytest_routed.txt
