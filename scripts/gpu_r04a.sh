#!/bin/bash
# final pass: the no-alloc entry point's test alone in a fresh process, the C++ routed-shards driver, then the whole GPU suite
# (which ends with the no-alloc test again) and smoke().  Outputs under gpurun_out/$TAG.
TAG=${1:-r04a}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?" | tee -a $O/rc.txt
tail -15 $O/pytest_no_alloc_alone.log
timeout 120 ./hellovulkan_b200/lib/hv_routed_shards_app 16 9 24 2560 1600 30000 12345 3 3 > $O/routed_shards_app.log 2>&1; echo "app rc=$?" | tee -a $O/rc.txt
cat $O/routed_shards_app.log | tail -5
timeout 600 python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -14 $O/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
tail -2 $O/smoke.log
cat $O/rc.txt
