"""Multi-GPU parity check, one rank per GPU (torchrun): the slab ring over N GPUs must reproduce the
single-GPU lattice bit for bit, for both transports (peer-to-peer stores with device flags; staged NCCL).
The cases live in latticeboltzmann_b200/ringcheck.py (bench.py --gpus N runs the same pass before timing).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py

--same-gpu: every rank uses GPU 0 and the handles travel over gloo -- on a one-GPU box this still exercises what a logical-slab
test cannot: cudaIpcOpenMemHandle peer mappings, the device-flag barrier between PROCESSES and the fused seam launches.
--abort: additionally, one rank stops stepping and the other must get an error after one time-out instead of hanging.
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200.ringcheck import abort_check, ring_parity  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    same = "--same-gpu" in sys.argv
    if same:
        local = 0
        torch.cuda.set_device(0)
        dist.init_process_group("gloo")
    else:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ok, total = ring_parity(local, log=lambda s: print(s, flush=True), staged=not same)
    if rank == 0:
        print(f"[mgpu] {ok}/{total} cases BIT-EXACT over {world} {'processes on one GPU' if same else 'GPUs'}", flush=True)
    bad = ok != total
    if "--abort" in sys.argv and world == 2:
        res = abort_check(local)
        if rank == 0:
            secs, text = res
            fine = "timed out" in text and secs < 15.0
            print(f"[mgpu] abort: rank 0 got '{text}' after {secs:.2f} s -> {'OK' if fine else 'FAILED'}", flush=True)
            bad = bad or not fine
    flag = [bad]
    dist.broadcast_object_list(flag, 0)
    dist.destroy_process_group()
    sys.exit(1 if flag[0] else 0)


if __name__ == "__main__":
    main()
