#!/bin/bash
# round 3, batch 6 (2 GPUs): final state -- suite, smoke, the driver's bench commands at N = 1 and N = 2
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_final_2gpu.log 2>&1; echo "suite rc=$?"
tail -4 gpurun_out/r03_pytest_final_2gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r03_bench_final_n1.json 2> gpurun_out/r03_bench_final_n1.err; echo "bench n1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29614 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r03_bench_final_n2.json 2> gpurun_out/r03_bench_final_n2.err; echo "bench n2 rc=$?"
python - <<PY
import json
for f in ("n1", "n2"):
    try:
        d=json.loads(open(f'gpurun_out/r03_bench_final_{f}.json').read().strip().split('\n')[-1])
        print(f, d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['summary'], d['errors'])
    except Exception as e:
        print(f, 'no line', e)
PY
