"""One multi-GPU block of bench.py on its own (under torchrun for N > 1), for iterating on it without the whole bench:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \\
        tools/run_block.py oplog_exchange            # or: mesh_route
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench_blocks as blocks  # noqa: E402
from radixmesh_b200 import capi  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "oplog_exchange"
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fn = {"oplog_exchange": blocks.oplog_exchange_block, "mesh_route": blocks.mesh_route_block}[which]
    out = fn(torch, capi, rank, world)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
