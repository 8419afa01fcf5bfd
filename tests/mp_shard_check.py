"""torchrun entry: one process per GPU.  Every rank runs the sharded SV sweep
collectively and then the same sweep unsharded on its own GPU; the shard of the
final particles and log Z must be bit-identical.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/mp_shard_check.py [n_local] [T]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    from anglican_b200 import Population
    from anglican_b200.distributed import nccl_unique_id
    from anglican_b200.models import StochasticVolatility
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    n_local = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = nccl_unique_id(rank)
    rng = np.random.default_rng(5)
    y = rng.normal(size=T) * 0.7
    m = StochasticVolatility()
    pop = Population(n_local, seed=2024, device=local, rank=rank, world=world, nccl_id=uid)
    ok = True
    for seed in (1, 2):
        pop.set_seed(seed)
        out = pop.sweep(m, y, want_filt=False)
        with Population(n_local * world, seed=seed, device=local) as ref_pop:
            ref = ref_pop.sweep(m, y, want_filt=False)
        shard = ref.x_final[rank * n_local:(rank + 1) * n_local]
        same = out.log_z == ref.log_z and np.array_equal(out.x_final, shard)
        print(f"rank {rank} seed {seed}: log_z {out.log_z!r} ref {ref.log_z!r} "
              f"bit-identical={same}", flush=True)
        ok = ok and same
    pop.close()
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    dist.destroy_process_group()
    if rank == 0:
        print("MP_SHARD_CHECK", "PASS" if int(flag.item()) == 0 else "FAIL", flush=True)
    sys.exit(0 if int(flag.item()) == 0 else 1)


if __name__ == "__main__":
    main()
