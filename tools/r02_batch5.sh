#!/bin/bash
set -x
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q ) > gpurun_out/b5_pytest.log 2>&1
tail -n 25 gpurun_out/b5_pytest.log
python tools/narrow_sweep.py > gpurun_out/b5_narrow.jsonl 2> gpurun_out/b5_narrow.err
python tools/gpu_sweep.py --kinds u64 --n 1000000000 --variants 0,4 --flags 1 --reps 3 > gpurun_out/b5_sweep_1B.jsonl 2> gpurun_out/b5_sweep_1B.err
python tools/gpu_sweep.py --kinds u32 --n 1000000000 --variants 0,5 --flags 1 --reps 3 >> gpurun_out/b5_sweep_1B.jsonl 2>> gpurun_out/b5_sweep_1B.err
