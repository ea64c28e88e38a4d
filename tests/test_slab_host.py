"""CPU: host-side logic of the slab decomposition -- row partition, ring neighbours, message
layout, and a world_size-2 (and 3) gloo run of the exchange plan where each rank steps its slab
with the CPU oracle; the stitched result must equal the single-slab oracle bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

from latticeboltzmann_b200 import slab

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_rows():
    assert slab.partition_rows(10, 3) == [(0, 4), (4, 7), (7, 10)]
    assert slab.partition_rows(8, 8) == [(k, k + 1) for k in range(8)]
    for nx, p in ((400, 7), (16384, 8), (5, 5)):
        b = slab.partition_rows(nx, p)
        assert b[0][0] == 0 and b[-1][1] == nx
        assert all(b[k][1] == b[k + 1][0] for k in range(p - 1))
        sizes = [hi - lo for lo, hi in b]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        slab.partition_rows(3, 4)


def test_ring_neighbours():
    assert slab.ring_neighbours(0, 4) == (1, 3)
    assert slab.ring_neighbours(3, 4) == (0, 2)
    assert slab.ring_neighbours(0, 2) == (1, 1)
    assert slab.ring_neighbours(0, 1) == (0, 0)


def test_message_layout():
    f = np.arange(9 * 5 * 4, dtype=np.float64).reshape(9, 5, 4)
    up, dn = slab.pack_messages(f)
    np.testing.assert_array_equal(up, f[[2, 5, 6], 3])
    np.testing.assert_array_equal(dn, f[[4, 7, 8], 1])
    g = np.zeros_like(f)
    slab.unpack_messages(g, up, dn)
    np.testing.assert_array_equal(g[[2, 5, 6], 0], up)
    np.testing.assert_array_equal(g[[4, 7, 8], 4], dn)
    assert g[[0, 1, 3, 4, 7, 8], 0].sum() == 0 and g[:, 1:4].sum() == 0


def _worker(rank, world, port, prec, nx, ny, steps, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from oracle.harness import Oracle
    from tests import cases
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    case = dict(prec=prec, nx=nx, ny=ny, scene="random", perturb=21)
    fA, walls = cases.make_inputs(case)
    lo, hi = slab.partition_rows(nx, world)[rank]
    nxl = hi - lo
    # extended slab: halo rows are marked solid so the oracle's ApplySource skips them and their own
    # (meaningless) update is a plain bounce; only rows 1..nxl are ever read back.
    o = Oracle(prec, nxl + 2, ny)
    ext_f = np.zeros((9, nxl + 2, ny), dtype=fA.dtype)
    ext_f[:, 1:-1] = fA[:, lo:hi]
    ext_w = np.ones((nxl + 2, ny), dtype=np.int8)
    ext_w[1:-1] = walls[lo:hi]
    o.set_state(ext_f, ext_w)
    tdt = torch.float64 if prec == "dp" else torch.float32
    for t in range(steps):
        src = "A" if t % 2 == 0 else "B"
        o.apply_source(src)                                   # Source on owned rows, then exchange
        view = o.f_view(src)
        to_up, to_dn = (torch.from_numpy(m) for m in slab.pack_messages(view))
        from_dn, from_up = torch.empty_like(to_up), torch.empty_like(to_dn)
        slab.exchange_halos(to_up, to_dn, from_dn, from_up, rank, world)
        slab.unpack_messages(view, from_dn.numpy(), from_up.numpy())
        o.stream_collide(src)
    np.save(os.path.join(out_dir, f"slab{rank}.npy"), o.f_view("A")[:, 1:-1])
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,prec", [(2, "dp"), (2, "sp"), (3, "dp")])
def test_gloo_ring_exchange_matches_single_slab(tmp_path, world, prec):
    import torch.multiprocessing as mp
    from oracle.harness import Oracle
    from tests import cases
    nx, ny, steps = 22, 48, 12
    mp.spawn(_worker, args=(world, _free_port(), prec, nx, ny, steps, str(tmp_path)), nprocs=world, join=True)
    fA, walls = cases.make_inputs(dict(prec=prec, nx=nx, ny=ny, scene="random", perturb=21))
    o = Oracle(prec, nx, ny)
    o.set_state(fA, walls)
    o.run(steps)
    got = np.concatenate([np.load(tmp_path / f"slab{r}.npy") for r in range(world)], axis=1)
    # wall-site f0 is never written: the single-slab run keeps the input there, the slab run too
    assert got.tobytes() == o.f_view("A").tobytes()
