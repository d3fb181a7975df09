#!/bin/bash
N=${1:-8}; TAG=${2:-r02f}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 600 python -m pytest tests/test_gpu_routed.py -x -q > $O/pytest_routed.txt 2>&1; echo "routed rc=$?"; tail -3 $O/pytest_routed.txt
bash scripts/gpu_r02d.sh $N $TAG "$3"
