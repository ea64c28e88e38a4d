"""Multi-GPU parity check, one rank per GPU (torchrun): the slab ring over N GPUs must reproduce the
single-GPU lattice bit for bit, for both transports (peer-to-peer stores with device flags; staged NCCL).
The cases live in latticeboltzmann_b200/ringcheck.py (bench.py --gpus N runs the same pass before timing).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200.ringcheck import ring_parity  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ok, total = ring_parity(local, log=lambda s: print(s, flush=True))
    if rank == 0:
        print(f"[mgpu] {ok}/{total} cases BIT-EXACT over {world} GPUs", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok == total else 1)


if __name__ == "__main__":
    main()
