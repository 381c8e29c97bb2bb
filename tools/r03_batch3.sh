#!/bin/bash
# round 3, batch 3 (8 GPUs): multi-GPU pytest on 8 devices; the driver's bench command at N = 8 and 4; the claim form at N = 8
mkdir -p gpurun_out
python -m pytest tests/test_mgpu_gpu.py tests/test_parity_gpu.py -x -q -m gpu -k "mgpu or communicator or every_device" > gpurun_out/r03_mgpu_pytest_8gpu.log 2>&1; echo "mgpu tests rc=$?"
tail -4 gpurun_out/r03_mgpu_pytest_8gpu.log
run() {  # name, nproc, extra args
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $2 --steps 10 --warmup 3 $3 > gpurun_out/r03_bench_$1.json 2> gpurun_out/r03_bench_$1.err; echo "bench $1 rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r03_bench_$1.json').read().strip().split('\n')[-1])
    print(d['value'], d['ms_per_step'], d['roofline']['phases_ms_rank0'], d['summary'], d['errors'])
except Exception as e:
    print('no line', e)
PY
  grep -n "Traceback\|Error\|assert" gpurun_out/r03_bench_$1.err | head -10
}
run n8 8 ""
run n8_claim 8 "--claim on --no-extra"
run n4 4 ""
