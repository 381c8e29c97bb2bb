#!/bin/bash
set -x
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/b4_pytest.log 2>&1
tail -n 15 gpurun_out/b4_pytest.log
python tools/narrow_sweep.py > gpurun_out/b4_narrow.jsonl 2> gpurun_out/b4_narrow.err
tail -n 3 gpurun_out/b4_narrow.err
python tools/gpu_sweep.py --kinds u64 --n 1000000000 --variants 0,3,4 --flags 1 --reps 3 > gpurun_out/b4_sweep_1B.jsonl 2> gpurun_out/b4_sweep_1B.err
python tools/gpu_sweep.py --kinds u32 --n 1000000000 --variants 0,2,4 --flags 1 --reps 3 >> gpurun_out/b4_sweep_1B.jsonl 2>> gpurun_out/b4_sweep_1B.err
