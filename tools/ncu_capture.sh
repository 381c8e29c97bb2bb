#!/bin/bash
# usage: tools/ncu_capture.sh <tag> <kernel-regex> <count> -- <python args for tools/one_sort.py>
# Runs the command plainly first, then under ncu --set full; exports the raw and source
# pages as CSV into gpurun_out/ and drops the (large) .ncu-rep unless it is small.
tag=$1; regex=$2; count=$3; shift 4
python tools/one_sort.py "$@" > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed"; tail -n 20 gpurun_out/plain_$tag.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$regex" -c $count -o /tmp/prof_$tag python tools/one_sort.py "$@" > gpurun_out/ncu_$tag.log 2>&1
ncu -i /tmp/prof_$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i /tmp/prof_$tag.ncu-rep --page source --csv > gpurun_out/${tag}_source.csv 2>/dev/null
ls -la /tmp/prof_$tag.ncu-rep gpurun_out/${tag}_*.csv
sz=$(stat -c %s /tmp/prof_$tag.ncu-rep)
if [ "$sz" -lt 30000000 ]; then cp /tmp/prof_$tag.ncu-rep gpurun_out/; fi
tail -n 2 gpurun_out/ncu_$tag.log
