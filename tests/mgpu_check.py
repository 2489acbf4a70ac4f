"""Run under torchrun (one rank per GPU): the sharded LMS sort + rank-0 induction must give the same suffix
array as the single-GPU path and as the oracle.  Used by tests/test_gpu_multi.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import kiss_b200  # noqa: E402
from kiss_b200.multi_gpu import sharded_suffix_array_dev  # noqa: E402
import oracle_lib as ol  # noqa: E402
import textgen as tg  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank = dist.get_rank()
    ctx = kiss_b200.Context(local)
    ok = True
    cases = [tg.uniform(3, 2_000_000), tg.planted_repeats(5, 1_500_000, unit_len=4000, copies=80),
             tg.mixed(9, 400_000), tg.uniform(4, 3000)]
    for ci, S in enumerate(cases):
        n = S.size
        dS = torch.zeros(n + 64, dtype=torch.uint8, device="cuda")
        dS[:n] = torch.from_numpy(S).cuda()
        SA = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        st = sharded_suffix_array_dev(ctx, dS, n, SA, 256, dist)
        if rank == 0:
            got = SA.cpu().numpy().view(np.uint32)
            want = ol.oracle_kiss1(S, k=256)
            same = np.array_equal(got, want)
            print(f"case {ci}: n={n} shards={st['shard_sizes']} equal={same}", flush=True)
            ok = ok and same
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
