#!/bin/bash
# Round evidence for profiles/: (1) plain bench run, (2) launch list of the same command
# (gpu__time_duration per launch), (3) DRAM traffic of one K3 launch at the bench size.
set -x
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/bench_plain.log 2>&1 || { tail -20 gpurun_out/bench_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_launches.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:onesweep -s 10 -c 2 --csv \
    --log-file gpurun_out/k3_traffic.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_traffic.log 2>&1
tail -3 gpurun_out/k3_traffic.csv
