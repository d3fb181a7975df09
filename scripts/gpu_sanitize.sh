#!/bin/bash
TAG=${1:-r03s}; O=gpurun_out/$TAG; mkdir -p $O
export PYTHONPATH=$PWD
for tool in memcheck racecheck synccheck; do
  for what in single rounds routed; do
    timeout 420 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_probe.py $what > $O/${tool}_$what.txt 2>&1
    echo "$tool $what rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|ok [0-9]' $O/${tool}_$what.txt | tr '\n' ' ')"
  done
done
