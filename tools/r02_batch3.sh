#!/bin/bash
set -x
mkdir -p gpurun_out
( time python bench.py --steps 5 --warmup 3 ) > gpurun_out/b3_bench_n1.json 2> gpurun_out/b3_bench_n1.err
tail -c 600 gpurun_out/b3_bench_n1.err
( time python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/b3_ref_n1.json 2> gpurun_out/b3_ref_n1.err
tail -c 300 gpurun_out/b3_ref_n1.err
free -g | head -2; nproc
