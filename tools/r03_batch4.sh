#!/bin/bash
# round 3, batch 4 (1 GPU): evidence.  Each ncu capture runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
# (0) the suite with this batch's additions (pooled workspace entry point, configs[2] at 1 B keys)
python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_1gpu.log 2>&1; echo "suite rc=$?"
tail -6 gpurun_out/r03_pytest_1gpu.log
# (1) plain default bench: the line, its summary, errors
python bench.py > gpurun_out/r03_bench_n1.json 2> gpurun_out/r03_bench_n1.err; echo "bench n1 rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r03_bench_n1.json').read().strip().split('\n')[-1])
print(d['value'], d['ms_per_step'], d['e2e'], d['roofline']['frac'], d['summary'], d['errors'])
PY
# (2) launch list + DRAM traffic of one K3 launch at the bench size
python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-e2e > gpurun_out/bench_plain.log 2>&1 || { tail -20 gpurun_out/bench_plain.log; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r03_bench_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-e2e > gpurun_out/ncu_launches.log 2>&1
python tools/launch_summary.py gpurun_out/r03_bench_launches.csv > gpurun_out/r03_bench_launch_list_summary.txt 2>&1; head -12 gpurun_out/r03_bench_launch_list_summary.txt
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:onesweep -s 10 -c 2 --csv \
    --log-file gpurun_out/r03_k3_traffic_1B_u64.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-e2e --no-u32 > gpurun_out/ncu_traffic.log 2>&1
tail -4 gpurun_out/r03_k3_traffic_1B_u64.csv
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:onesweep -s 2 -c 2 --csv \
    --log-file gpurun_out/r03_k3_traffic_1B_u32.csv python tools/one_sort.py --kind u32 --n 1000000000 --reps 1 > gpurun_out/ncu_traffic32.log 2>&1
tail -3 gpurun_out/r03_k3_traffic_1B_u32.csv
# (3) ncu --set full digests: K3 u64 / u32 at 2^27 (middle passes), the counting path of 1- and 2-byte keys at 2^30
bash tools/ncu_capture.sh r03_k3_u64_2p27 onesweep 4 -- --kind u64 --n 134217728 --reps 1
bash tools/ncu_capture.sh r03_k3_u32_2p27 onesweep 3 -- --kind u32 --n 134217728 --reps 1
bash tools/ncu_capture.sh r03_narrow_u8_2p30 "hist_values|fill_values|scan_values" 3 -- --kind u8 --n 1073741824 --reps 1
bash tools/ncu_capture.sh r03_narrow_u16_2p30 "hist_values|fill_values|reduce_partials|scan_values" 4 -- --kind u16 --n 1073741824 --reps 1
for t in r03_k3_u64_2p27 r03_k3_u32_2p27 r03_narrow_u8_2p30 r03_narrow_u16_2p30; do python tools/ncu_raw.py gpurun_out/${t}_raw.csv > gpurun_out/${t}.txt 2>&1; done
rm -f gpurun_out/*_source.csv gpurun_out/*.ncu-rep
ls -la gpurun_out | tail -30
