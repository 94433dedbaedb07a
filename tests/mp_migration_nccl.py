"""torchrun script (one process per GPU, NCCL): displacement + migration between slabs, checked
record by record against the host restatement.  Launched by tests/test_gpu_migration.py when
the box has >= 2 GPUs:  python -m torch.distributed.run --nproc-per-node 2 tests/mp_migration_nccl.py"""

import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tests.migration_common import Shard, global_population, sort_and_count  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    try:
        pop = global_population()
        shard = Shard(pop, rank, world)
        shard.displace()
        shard.migrator.migrate()  # pack -> NCCL all_to_all_single (counts, records) -> unpack
        n = shard.check(pop)
        sizes = sort_and_count(shard)
        total = torch.tensor([n, shard.migrator.n_sent, shard.migrator.n_received],
                             dtype=torch.int64, device="cuda")
        dist.all_reduce(total)
        assert int(total[0]) == pop["multiplicity"].size and int(total[1]) == int(total[2]) > 0
        assert sizes.sum() == n
        print(f"migration ok rank {rank}: holds {n}, sent {shard.migrator.n_sent}, "
              f"received {shard.migrator.n_received}", flush=True)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
