"""Round 2, torchrun: weak-scaling unit per GPU over N GPUs, MLUPS over strip width / hints / concurrent chains of the slab schedule
(the seam chain costs one handshake per strip and level).  usage (torchrun): mgpu_sweep.py [dp|sp|both]"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200.slab import DistributedLattice  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
precs = ("dp", "sp") if len(sys.argv) < 2 or sys.argv[1] == "both" else (sys.argv[1],)
nx, ny, T = 8192 * world, 32768, 96


def mx(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for prec in precs:
    R = 8 if prec == "dp" else 4
    dl = DistributedLattice(nx, ny, prec, device=local, inplace=True)
    dl.lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=min(nx, ny) // 10)
    combos = [(16, 0, 0, 1), (16, 0, 0, 2), (16, 0, 3, 2), (16, 16384 // R, 0, 1), (16, 16384 // R, 0, 2), (16, 16384 // R, 3, 2),
              (16, 0, 0, 3), (16, 16384 // R, 0, 3), (16, 65536 // R, 0, 2)]
    for k, w, hints, chains in combos:
        dl.lat.set_tuning(tb_depth=k, tb_cols=w, hints=hints, tb_chains=chains)
        dist.barrier()
        dl.run_timed(k)
        best = 1e9
        for _ in range(2):
            torch.cuda.synchronize(); dist.barrier()
            best = min(best, mx(dl.run_timed(T)))
        if rank == 0:
            print(json.dumps(dict(prec=prec, gpus=world, k=k, W=w, hints=hints, chains=chains, mlups=round(nx * ny * T / best / 1e3, 1),
                                  per_gpu=round(nx * ny * T / best / 1e3 / world, 1))), flush=True)
    dl.close()
    dist.barrier()
dist.destroy_process_group()
