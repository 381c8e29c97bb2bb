"""2-GPU probe (torchrun): is torch symmetric memory usable here, and does the
fused K7 exchange give the same result as the all-to-all path?  Dev tool."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import radix_sort_b200 as rs
from radix_sort_b200 import distr, mgpu, verify

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
K = rs.Kind.U64
src = distr.empty_keys(n, K, dev)
distr.generate(src, K, distr.Distr.UNIFORM, 0, seed=7, index_base=rank * n)
ws = torch.empty(rs.temp_bytes(int(n * 1.25) + 1024, K), dtype=torch.uint8, device=dev)
t0 = time.time()
ex = mgpu.PeerExchange(int(n * 1.25) + 1024, 8, dev)
torch.cuda.synchronize()
print(f"rank {rank}: PeerExchange ok in {time.time() - t0:.2f}s ptrs={[hex(p) for p in ex.peer_ptrs]}", flush=True)
for trial in range(3):
    keys = src.clone()
    ph = {}
    a = mgpu.sort_sharded(keys, K, workspace=ws, timings=ph)
    ref = a.keys.clone()
    keys = src.clone()
    ph2 = {}
    b = mgpu.sort_sharded(keys, K, workspace=ws, exchange=ex, timings=ph2)
    torch.cuda.synchronize()
    same = a.n_out == b.n_out and bool((ref.view(torch.int64) == b.keys.view(torch.int64)).all())
    print(f"rank {rank} trial {trial}: n_out {a.n_out} {b.n_out} identical={same} unsorted={verify.count_unsorted(b.keys, K)}\n"
          f"   nccl  phases {dict((k, round(v, 2)) for k, v in ph.items())}\n"
          f"   fused phases {dict((k, round(v, 2)) for k, v in ph2.items())}", flush=True)
    assert same
dist.barrier()
dist.destroy_process_group()
