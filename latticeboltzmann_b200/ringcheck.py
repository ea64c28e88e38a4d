"""Ring parity over real GPUs, one rank per GPU (torchrun): the slab ring -- peer stores over NVLink / CUDA IPC, device-flag
ordering, temporal blocking across the seams -- must reproduce the single-GPU lattice bit for bit.

Used by tools/mgpu_check.py, tests/test_multi_gpu.py and by ``bench.py --gpus N`` (N >= 2), which runs it before timing and
stamps the outcome into ``config.parity``.  The single-GPU side is the A/B engine of the same library, itself pinned
against the compiled reference's golden vectors by tests/test_gpu_parity.py.
"""
import time

import numpy as np

from . import geometry


def ring_cases(world):
    """(prec, transport, nx, ny, engine): engine False = A/B, True = in place, dict = in place + that tuning."""
    return [("dp", "p2p", 803, 2000, False), ("sp", "p2p", 803, 2000, False), ("dp", "staged", 401, 1000, False),
            ("dp", "p2p", world, 64, False), ("sp", "p2p", 2 * world + 1, 264, False),
            # one buffer per slab, in-place kernels: phase-1 boundary rows store into the neighbour's OWNED rows
            ("dp", "p2p", 803, 2000, True), ("sp", "p2p", 803, 2000, True),
            ("dp", "p2p", world, 64, True), ("sp", "p2p", 2 * world + 1, 264, True),
            # + temporal blocking across the slab seams: (depth, band rows, strip columns); slabs too low for the
            # row pyramid take whole-slab launches per event
            ("dp", "p2p", 803, 2000, dict(tb_depth=4, tb_rows=50, tb_cols=0)), ("sp", "p2p", 803, 2000, dict(tb_depth=6, tb_rows=0, tb_cols=256)),
            ("dp", "p2p", 803, 2000, dict(tb_depth=16, tb_rows=33, tb_cols=512, vec=1)), ("dp", "p2p", world, 64, dict(tb_depth=2, tb_rows=0, tb_cols=0)),
            ("sp", "p2p", 2 * world + 1, 264, dict(tb_depth=4, tb_rows=0, tb_cols=64)), ("dp", "p2p", 7 * world + 3, 520, dict(tb_depth=4, tb_rows=2, tb_cols=104)),
            # + concurrent tile chains inside every slab (lanes on their own streams; the slab seam is the last lane boundary)
            ("dp", "p2p", 803, 2000, dict(tb_depth=4, tb_rows=20, tb_cols=0, tb_chains=2)),
            ("sp", "p2p", 803, 2000, dict(tb_depth=6, tb_rows=0, tb_cols=256, tb_chains=3))]


def ring_inputs(prec, nx, ny):
    fA = geometry.equilibrium_rest(nx, ny, prec)
    if nx > 100:
        return fA, geometry.cylinder_walls(nx, ny, nx // 2, ny // 5, nx // 10)
    fA = geometry.perturb_lcg(fA, seed=3)
    walls = np.zeros((nx, ny), dtype=np.int8)                # tiny lattices: no channel rows (they would leave no fluid row)
    walls[0, ::5] = 1
    return fA, walls


def ring_parity(local_device, log=None, staged=True):
    """Runs every case of ``ring_cases`` over the ranks of the default process group; returns (passed, total) on every rank.
    ``staged=False`` leaves out the NCCL-staged transport (ranks sharing one GPU talk over gloo)."""
    import torch.distributed as dist
    from .slab import DistributedLattice
    from .solver import Lattice
    rank, world = dist.get_rank(), dist.get_world_size()
    cases = [c for c in ring_cases(world) if staged or c[1] != "staged"]
    nsame = 0
    for prec, transport, nx, ny, engine in cases:
        fA, walls = ring_inputs(prec, nx, ny)
        dl = DistributedLattice(nx, ny, prec, transport, device=local_device, inplace=bool(engine))
        if isinstance(engine, dict):
            dl.lat.set_tuning(**engine)
        dl.lat.upload(fA, walls)
        dist.barrier()
        t0 = time.time()
        dl.run(20)
        dl.run(40)
        mine = dl.lat.download_slab()
        mine_b = None if engine else dl.lat.download_slab(which="B")
        mine_sum = dl.lat.checksum()
        dt = time.time() - t0
        re = dl.gather_reynolds()
        parts, parts_b, sums = [None] * world, [None] * world, [None] * world
        dist.all_gather_object(parts, mine)
        dist.all_gather_object(parts_b, mine_b)
        dist.all_gather_object(sums, mine_sum)
        if rank == 0:
            with Lattice(nx, ny, prec, device=local_device) as one:
                one.upload(fA, walls)
                one.run(60)
                want, want_b, want_re, want_sum = one.download("A"), one.download("B"), one.ComputeReynolds(), one.checksum()
            got = np.concatenate(parts, axis=1)
            same = got.tobytes() == want.tobytes() and float(re).hex() == float(want_re).hex() and sum(sums) % 2 ** 64 == want_sum
            if not engine:
                same = same and np.concatenate(parts_b, axis=1).tobytes() == want_b.tobytes()
            nsame += int(same)
            if log:
                tag = " in-place" if engine else ""
                tag += " " + str(engine) if isinstance(engine, dict) else ""
                log(f"[mgpu] {prec} {transport}{tag} {nx}x{ny} over {world} GPUs: {'BIT-EXACT' if same else 'MISMATCH'} vs single GPU (A/B) "
                    f"(Re {re!r} vs {want_re!r}; {dt:.2f} s)")
                if not same:
                    log("   first differing elements (s,i,j): " + str(np.argwhere(got != want)[:5].tolist()))
        dl.close()
        dist.barrier()
    box = [nsame]
    dist.broadcast_object_list(box, 0)                       # (works under NCCL and under gloo)
    return int(box[0]), len(cases)


def abort_check(local_device, timeout_ms=400):
    """One rank of a 2-rank ring stops stepping: the other rank's waits must give up after ONE time-out (the ring's abort word),
    lbm_sync must report it, and nothing may hang.  Returns (seconds rank 0 spent, error text) on rank 0."""
    import os
    import torch.distributed as dist
    from ._capi import LbmError
    from .slab import DistributedLattice
    rank = dist.get_rank()
    os.environ["LBM_WAIT_TIMEOUT_MS"] = str(timeout_ms)
    try:
        nx, ny = 64, 264
        fA, walls = ring_inputs("dp", nx, ny)
        dl = DistributedLattice(nx, ny, "dp", "p2p", device=local_device, inplace=True)
        dl.lat.set_tuning(tb_depth=4, tb_rows=4, tb_cols=64)          # dozens of seam waits per run
        dl.lat.upload(fA, walls)
        dist.barrier()
        dl.run(8)
        dl.lat.sync()
        dist.barrier()
        out = None
        if rank == 0:                                            # rank 1 does not step any more
            t0 = time.time()
            try:
                dl.run(40)
                dl.lat.sync()
                out = (time.time() - t0, "no error")
            except LbmError as exc:
                out = (time.time() - t0, str(exc))
        dist.barrier()
        dl.close()
        return out
    finally:
        os.environ.pop("LBM_WAIT_TIMEOUT_MS", None)
