"""Row-slab decomposition over ranks (one process per GPU) -- the B200 replacement of the
reference's MPI row slabs (README.md:44-53; the MPI source is not in the checkout, so the
exchange is derived from the stencil, LB.c:231-243).

Rank r owns rows [lo_r, hi_r).  A pull step needs, from outside the slab,
  * row lo-1 (wrapping to nx-1): speeds 2, 5, 6   (LB.c:237,240,241)
  * row hi   (wrapping to 0):    speeds 4, 7, 8   (LB.c:239,242,243)
so every step each rank sends its LAST row's f2,f5,f6 up (to rank r+1) and its FIRST row's
f4,f7,f8 down (to rank r-1), a periodic ring; 3*ny reals per direction.

Two transports:
  * ``p2p``   (default): CUDA-IPC peer mappings; the boundary-row kernel stores the halo speeds
    straight into the neighbour's halo rows over NVLink and steps are ordered by device flags.
    torch.distributed only ships the 256-byte handle blobs once.
  * ``staged``: pack -> torch.distributed isend/irecv (NCCL on GPUs, gloo in CPU tests) -> unpack.
"""
import numpy as np

HALO_UP_SPEEDS = (2, 5, 6)   # travel to the slab above (rank r+1)
HALO_DN_SPEEDS = (4, 7, 8)   # travel to the slab below (rank r-1)


def partition_rows(nx, nparts):
    """Contiguous row ranges, sizes differing by at most one (OpenMP schedule(static), LB.c:328)."""
    if nparts < 1 or nparts > nx:
        raise ValueError(f"cannot split {nx} rows into {nparts} slabs")
    base, extra = divmod(nx, nparts)
    bounds, lo = [], 0
    for k in range(nparts):
        hi = lo + base + (1 if k < extra else 0)
        bounds.append((lo, hi))
        lo = hi
    return bounds


def ring_neighbours(rank, world):
    """(up, dn): ranks owning the rows above row_hi and below row_lo, periodic (LB.c:231-232)."""
    return (rank + 1) % world, (rank - 1) % world


def pack_messages(f_ext):
    """Host statement of the message layout (mirrors halo_pack_kernel): ``f_ext`` is (9, nxl+2, ny) with
    halo rows 0 and nxl+1.  Returns (to_up, to_dn), each (3, ny): the last owned row's speeds 2,5,6 and
    the first owned row's speeds 4,7,8."""
    return (np.ascontiguousarray(f_ext[list(HALO_UP_SPEEDS), -2, :]),
            np.ascontiguousarray(f_ext[list(HALO_DN_SPEEDS), 1, :]))


def unpack_messages(f_ext, from_dn, from_up):
    """Mirrors halo_unpack_kernel: what the slab below sent up lands in the lower halo row (speeds
    2,5,6), what the slab above sent down lands in the upper halo row (speeds 4,7,8)."""
    f_ext[list(HALO_UP_SPEEDS), 0, :] = from_dn
    f_ext[list(HALO_DN_SPEEDS), -1, :] = from_up


def exchange_halos(to_up, to_dn, from_dn, from_up, rank, world, group=None):
    """One ring exchange of packed halo messages (torch tensors; CPU/gloo or CUDA/NCCL).

    ``to_up`` goes to rank+1 and arrives there as ``from_dn``; ``to_dn`` goes to rank-1 and
    arrives as ``from_up``.  world == 1 degenerates to the periodic self-wrap."""
    import torch.distributed as dist
    if world == 1:
        from_dn.copy_(to_up)
        from_up.copy_(to_dn)
        return
    up, dn = ring_neighbours(rank, world)
    ops = [dist.P2POp(dist.isend, to_up, up, group, tag=0), dist.P2POp(dist.irecv, from_dn, dn, group, tag=0),
           dist.P2POp(dist.isend, to_dn, dn, group, tag=1), dist.P2POp(dist.irecv, from_up, up, group, tag=1)]
    for w in dist.batch_isend_irecv(ops):
        w.wait()


def allgather_bytes(blob, world, group=None):
    """All-gather a fixed-size uint8 numpy blob over torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.from_numpy(np.ascontiguousarray(blob, dtype=np.uint8)).to(dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return [t.cpu().numpy() for t in out]


class DistributedLattice:
    """This rank's slab of a lattice spread over ``world`` ranks (one process per GPU)."""

    def __init__(self, nx, ny, prec="dp", transport="p2p", group=None, device=None, rank=None, world=None, **kw):
        import torch.distributed as dist
        from .solver import Lattice
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.group, self.transport = group, transport
        if kw.get("inplace") and transport == "staged" and self.world > 1:
            raise ValueError("the staged transport needs the A/B layout; in-place slabs exchange by peer stores (transport='p2p')")
        self.bounds = partition_rows(nx, self.world)
        lo, hi = self.bounds[self.rank]
        if device is None:
            import torch
            device = torch.cuda.current_device()
        self.lat = Lattice(nx, ny, prec, device=device, row_lo=lo, row_hi=hi, **kw)
        self.nx, self.ny = nx, ny
        if self.world > 1 and transport == "p2p":
            blobs = allgather_bytes(self.lat.ipc_export(), self.world, group)
            up, dn = ring_neighbours(self.rank, self.world)
            self.lat.ipc_connect(blobs[up], blobs[dn])
            dist.barrier(group)
        elif self.world > 1 and transport == "staged":
            import torch
            n = 3 * ny
            tdt = torch.float64 if prec == "dp" else torch.float32
            self._msg = [torch.zeros(n, dtype=tdt, device=f"cuda:{device}") for _ in range(4)]
            import ctypes
            sp = ctypes.c_void_p()
            self.lat._call("lbm_stream_handle", 1, ctypes.byref(sp))
            self._bstream = torch.cuda.ExternalStream(sp.value, device=f"cuda:{device}")

    # ---- stepping ---------------------------------------------------------------------------
    def run(self, nsteps):
        if self.world == 1 or self.transport == "p2p":
            self.lat.run(nsteps)
        else:
            self._run_staged(nsteps)

    def run_timed(self, nsteps):
        """Device time of this rank (CUDA events on the slab's stream), milliseconds."""
        if self.world == 1 or self.transport == "p2p":
            return self.lat.run_timed(nsteps)
        import torch
        self.lat.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self._bstream):
            e0.record()
            self._run_staged(nsteps)
            e1.record()
        self.lat.sync()
        return e0.elapsed_time(e1)

    def _run_staged(self, nsteps):
        """Boundary rows -> pack -> NCCL ring exchange -> unpack on the boundary stream while the
        inner rows run on the main stream (README.md:45-51)."""
        import ctypes
        import torch
        assert nsteps % 2 == 0
        lat, nxl = self.lat, self.lat.nx_local
        to_up, to_dn, from_dn, from_up = self._msg
        p = [ctypes.c_void_p(t.data_ptr()) for t in self._msg]

        def refresh(buf):
            lat._call("lbm_halo_pack", buf, p[0], p[1])
            with torch.cuda.stream(self._bstream):
                exchange_halos(to_up, to_dn, from_dn, from_up, self.rank, self.world, self.group)
            lat._call("lbm_halo_unpack", buf, p[2], p[3])

        # Source(A) of the first step, then make the neighbours' copies of our boundary rows current
        self._apply_source_staged()
        lat._call("lbm_stream_join")
        refresh(0)
        lat._call("lbm_stream_join")
        for t in range(nsteps):
            src, tail = t & 1, int(t != nsteps - 1)
            lat._call("lbm_step_rows", src, 0, 1, tail, 1)
            if nxl > 1:
                lat._call("lbm_step_rows", src, nxl - 1, nxl, tail, 1)
            lat._call("lbm_step_rows", src, 1, nxl - 1, tail, 0)
            refresh(src ^ 1)
            lat._call("lbm_stream_join")

    def _apply_source_staged(self):
        # lbm_apply_source refuses wired slabs; a staged slab is unwired (its own neighbour)
        self.lat.ApplySource("A")

    # ---- observers -----------------------------------------------------------------------------
    def gather_reynolds(self):
        """ComputeReynolds over the ring: centre-column u_y of every slab, concatenated in row order and
        summed serially on every rank (LB.c:529-546)."""
        import torch
        import torch.distributed as dist
        uy, fluid = self.lat.reynolds_column()
        if self.world == 1:
            return self.lat.reynolds_from_column(uy, fluid)
        objs = [None] * self.world
        dist.all_gather_object(objs, (uy, fluid), group=self.group)
        return self.lat.reynolds_from_column(np.concatenate([o[0] for o in objs]), np.concatenate([o[1] for o in objs]))

    def close(self):
        self.lat.close()
