#!/bin/bash
set -x
mkdir -p gpurun_out
python tools/gpu_sweep.py --kinds u32 --n 268435456 --variants 0,4,9,10,11,12 --check --flags 1 > gpurun_out/b2_sweep.jsonl 2> gpurun_out/b2_sweep.err
python tools/gpu_sweep.py --kinds u64 --n 268435456 --variants 0,3,8,9,10,11 --check --flags 1 >> gpurun_out/b2_sweep.jsonl 2>> gpurun_out/b2_sweep.err
cat gpurun_out/b2_sweep.jsonl | cut -c1-300
