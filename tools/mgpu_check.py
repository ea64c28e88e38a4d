"""Multi-GPU parity check, one rank per GPU (torchrun): the slab ring over N GPUs must reproduce the
single-GPU lattice bit for bit, for both transports (peer-to-peer stores with device flags; staged NCCL).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py
"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402
from latticeboltzmann_b200.slab import DistributedLattice  # noqa: E402
from tests import cases  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ok, nsame = True, 0
    CASES = [("dp", "p2p", 803, 2000, False), ("sp", "p2p", 803, 2000, False), ("dp", "staged", 401, 1000, False),
             ("dp", "p2p", world, 64, False), ("sp", "p2p", 2 * world + 1, 264, False),
             # one buffer per slab, in-place kernels: phase-1 boundary rows store into the neighbour's OWNED rows
             ("dp", "p2p", 803, 2000, True), ("sp", "p2p", 803, 2000, True),
             ("dp", "p2p", world, 64, True), ("sp", "p2p", 2 * world + 1, 264, True),
             # + temporal blocking across the slab seams: (depth, band rows, strip columns); slabs too low for the
             # row pyramid take whole-slab launches per event
             ("dp", "p2p", 803, 2000, dict(tb_depth=4, tb_rows=50, tb_cols=0)), ("sp", "p2p", 803, 2000, dict(tb_depth=6, tb_rows=0, tb_cols=256)),
             ("dp", "p2p", 803, 2000, dict(tb_depth=16, tb_rows=33, tb_cols=512, vec=1)), ("dp", "p2p", world, 64, dict(tb_depth=2, tb_rows=0, tb_cols=0)),
             ("sp", "p2p", 2 * world + 1, 264, dict(tb_depth=4, tb_rows=0, tb_cols=64)), ("dp", "p2p", 7 * world + 3, 520, dict(tb_depth=4, tb_rows=2, tb_cols=104))]
    for prec, transport, nx, ny, inplace in CASES:
        c = dict(prec=prec, isa="avx", nx=nx, ny=ny, scene="cylinder" if nx > 100 else "random", perturb=None if nx > 100 else 3)
        fA, walls = cases.make_inputs(c)
        if nx <= 100:                              # tiny lattices: no channel rows (they would leave no fluid row)
            walls[:] = 0
            walls[0, ::5] = 1
        dl = DistributedLattice(nx, ny, prec, transport, device=local, inplace=bool(inplace))
        if isinstance(inplace, dict):
            dl.lat.set_tuning(**inplace)
        dl.lat.upload(fA, walls)
        dist.barrier()
        t0 = time.time()
        dl.run(20)
        dl.run(40)
        mine = dl.lat.download_slab()
        mine_b = None if inplace else dl.lat.download_slab(which="B")
        dt = time.time() - t0
        re = dl.gather_reynolds()
        parts, parts_b = [None] * world, [None] * world
        dist.all_gather_object(parts, mine)
        dist.all_gather_object(parts_b, mine_b)
        if rank == 0:
            with Lattice(nx, ny, prec, device=local) as one:
                one.upload(fA, walls)
                one.run(60)
                want, want_b, want_re = one.download("A"), one.download("B"), one.ComputeReynolds()
            got = np.concatenate(parts, axis=1)
            same = got.tobytes() == want.tobytes() and float(re).hex() == float(want_re).hex()
            if not inplace:
                same = same and np.concatenate(parts_b, axis=1).tobytes() == want_b.tobytes()
            ok &= same
            nsame += int(same)
            print(f"[mgpu] {prec} {transport}{' in-place' if inplace else ''}{' ' + str(inplace) if isinstance(inplace, dict) else ''} {nx}x{ny} over {world} GPUs: {'BIT-EXACT' if same else 'MISMATCH'} vs single GPU (A/B) "
                  f"(Re {re!r} vs {want_re!r}; {dt:.2f} s)", flush=True)
            if not same:
                bad = np.argwhere(got != want)
                print("   first differing elements (s,i,j):", bad[:5].tolist(), flush=True)
        dl.close()
        dist.barrier()
    if rank == 0:
        print(f"[mgpu] {nsame}/{len(CASES)} cases BIT-EXACT over {world} GPUs", flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
