#!/bin/bash
# round 3, batch 1 (2 GPUs): the no-alloc entry point alone and after the suite; the 2-GPU bench arm with its stderr
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r03_zz_alone.log 2>&1; echo "zz alone rc=$?"
tail -3 gpurun_out/r03_zz_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_2gpu.log 2>&1; echo "suite rc=$?"
tail -5 gpurun_out/r03_pytest_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r03_bench_n2.json 2> gpurun_out/r03_bench_n2.err; echo "bench n2 rc=$?"
tail -c 1500 gpurun_out/r03_bench_n2.json
grep -v "^\s*$" gpurun_out/r03_bench_n2.err | grep -n "Error\|error\|Traceback\|File \"/root\|raise\|assert" | head -40
