#!/bin/bash
# round 3, batch 2 (2 GPUs): suite after the stagger removal + claim tests; 2-GPU bench with and without the claim form
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_2gpu_b.log 2>&1; echo "suite rc=$?"
tail -15 gpurun_out/r03_pytest_2gpu_b.log
for claim in off on; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 --claim $claim > gpurun_out/r03_bench_n2_claim_$claim.json 2> gpurun_out/r03_bench_n2_claim_$claim.err; echo "bench n2 claim=$claim rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r03_bench_n2_claim_$claim.json').read().strip().split('\n')[-1])
    print(d['value'], d['ms_per_step'], d['roofline']['phases_ms_rank0'], d['summary'], d['errors'])
except Exception as e:
    print('no line', e)
PY
grep -n "Traceback\|Error\|assert" gpurun_out/r03_bench_n2_claim_$claim.err | head -10
done
