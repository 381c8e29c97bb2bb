#!/bin/bash
# round 3, batch 5 (1 GPU): word-level narrow histogram -- suite, sweep, ncu digests after the change
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_1gpu_b5.log 2>&1; echo "suite rc=$?"
tail -4 gpurun_out/r03_pytest_1gpu_b5.log
python tools/narrow_sweep.py > gpurun_out/r03_narrow_keys_word_histogram.jsonl 2> gpurun_out/narrow_sweep.err; echo "sweep rc=$?"
grep '"histogram+fill"' gpurun_out/r03_narrow_keys_word_histogram.jsonl | grep -v '"n": 16777216' | cut -c1-200
bash tools/ncu_capture.sh r03_narrow_u8_2p30_word "hist_values|fill_values" 2 -- --kind u8 --n 1073741824 --reps 1
bash tools/ncu_capture.sh r03_narrow_u16_2p30_word "hist_values|fill_values" 2 -- --kind u16 --n 1073741824 --reps 1
for t in r03_narrow_u8_2p30_word r03_narrow_u16_2p30_word; do python tools/ncu_raw.py gpurun_out/${t}_raw.csv > gpurun_out/${t}.txt 2>&1; grep -E "=====|time_duration|stalls|lsu_wavefronts.avg|dram_throughput" gpurun_out/${t}.txt; done
rm -f gpurun_out/*_source.csv gpurun_out/*.ncu-rep
# where should the bounce-buffer path of b200sort_keys_host start?  C1's host call (10 M u32 = 40 MB pageable) by threshold
for thr in 67108864 33554432 16777216 4194304; do
  echo "{\"stage_min_bytes\": $thr}" >> gpurun_out/r03_host_latency_stage_min.jsonl
  B200SORT_STAGE_MIN_BYTES=$thr python tools/host_latency.py 2>/dev/null | grep -E '"n": (1000000|10000000)\b' >> gpurun_out/r03_host_latency_stage_min.jsonl
done
cat gpurun_out/r03_host_latency_stage_min.jsonl | cut -c1-120
