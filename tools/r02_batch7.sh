#!/bin/bash
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "narrow or counting or golden or edge" ) > gpurun_out/b7_pytest.log 2>&1
tail -n 5 gpurun_out/b7_pytest.log
python tools/narrow_sweep.py > gpurun_out/b7_narrow.jsonl 2> gpurun_out/b7_narrow.err
