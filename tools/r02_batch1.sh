#!/bin/bash
# round-2 GPU batch 1: sanity (pytest -m gpu), K3 variant sweep, distribution families, stagger experiment
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/b1_smi.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/b1_pytest.log 2>&1
tail -n 5 gpurun_out/b1_pytest.log
python tools/gpu_sweep.py --kinds u32,u64 --n 268435456 --variants all --check --flags 1 > gpurun_out/b1_sweep_uniform.jsonl 2> gpurun_out/b1_sweep_uniform.err
for d in "ZIPF 1.0" "ASCENDING 0" "DESCENDING 0" "STEP 16" "GEOMETRIC 0.0009765625"; do
  set -- $d
  python tools/gpu_sweep.py --kinds u32 --n 268435456 --variants 0,5,6 --check --distr $1 --param $2 >> gpurun_out/b1_sweep_families.jsonl 2>> gpurun_out/b1_sweep_families.err
  python tools/gpu_sweep.py --kinds u64 --n 268435456 --variants 0,4,5 --check --distr $1 --param $2 >> gpurun_out/b1_sweep_families.jsonl 2>> gpurun_out/b1_sweep_families.err
done
for ns in 0 1500 3000 5000 8000; do
  echo "stagger_ns=$ns" >> gpurun_out/b1_stagger.jsonl
  B200SORT_STAGGER_NS=$ns python tools/gpu_sweep.py --kinds u32,u64 --n 268435456 --variants 0 --flags 1 >> gpurun_out/b1_stagger.jsonl 2>> gpurun_out/b1_stagger.err
done
# 1 B keys, the best candidates
python tools/gpu_sweep.py --kinds u64 --n 1000000000 --variants 0,5,6 --flags 1 --reps 2 > gpurun_out/b1_sweep_1B.jsonl 2> gpurun_out/b1_sweep_1B.err
python tools/gpu_sweep.py --kinds u32 --n 1000000000 --variants 0,6,7 --flags 1 --reps 2 >> gpurun_out/b1_sweep_1B.jsonl 2>> gpurun_out/b1_sweep_1B.err
echo done
