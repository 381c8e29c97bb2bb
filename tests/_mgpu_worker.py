"""Worker of tests/test_mgpu_gpu.py: one rank of a 2-GPU sharded sort (torchrun).
Checks the fused peer-store exchange and the NCCL all-to-all path against numpy."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import radix_sort_b200 as rs  # noqa: E402
from radix_sort_b200 import mgpu  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        exchange = mgpu.PeerExchange(400_000, 8, dev)
        for case, n in (("uniform", 250_000), ("few_unique", 250_000), ("ragged", 250_000 if rank == 0 else 0),
                        ("overflow", 500_000)):  # overflow: shards exceed the exchange capacity -> fallback path
            rng = np.random.default_rng(77 + rank)
            a = rng.integers(0, 2**64, size=n, dtype=np.uint64)
            if case == "few_unique":
                a = (a % 5) * 1000
            for use_exchange, overlap in ((True, False), (True, True), (False, False)):
                keys = torch.from_numpy(a.copy()).to(dev)
                res = mgpu.sort_sharded(keys, rs.Kind.U64, exchange=exchange if use_exchange else None, overlap=overlap)
                mine = res.keys.cpu().numpy().copy()
                gathered = [None] * world
                dist.all_gather_object(gathered, (a, mine))
                whole_in = np.concatenate([g[0] for g in gathered])
                whole_out = np.concatenate([g[1] for g in gathered])
                assert np.array_equal(whole_out, np.sort(whole_in)), (case, use_exchange, overlap)
                torch.cuda.synchronize()
                dist.barrier()
        if rank == 0:
            print("MGPU_OK", flush=True)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
