"""CPU: the oracle's restatement of the boundary options the reference names but does not implement (open y ends,
LB.c:42-43; reflect walls, LB.c:21) -- parity unpinned, so what can be checked here are the definitions' own properties."""
import numpy as np

from latticeboltzmann_b200 import geometry
from oracle.harness import Oracle


def _run(nx, ny, walls, steps, prec="dp", **kw):
    o = Oracle(prec, nx, ny, **kw)
    o.set_state(geometry.perturb_lcg(geometry.equilibrium_rest(nx, ny, prec), seed=5), walls)
    o.run(steps)
    return o


def test_reflect_walls_conserve_mass_and_slip():
    nx, ny = 24, 96
    noslip = geometry.channel_walls(nx, ny, barrier=False)
    slip = noslip * 2
    a, b = _run(nx, ny, noslip, 300), _run(nx, ny, slip, 300)
    m0 = geometry.perturb_lcg(geometry.equilibrium_rest(nx, ny, "dp"), seed=5).astype(np.float64).sum()
    for o in (a, b):                                       # a permutation of a wall site's populations conserves mass
        assert abs(o.f_view("A").astype(np.float64).sum() - m0) < 1e-9 * m0
    ua, ub = a.macroscopic()[2], b.macroscopic()[2]
    assert ub[1:-1].mean() > 1.2 * ua[1:-1].mean() > 0     # no wall friction: the same body force drives a faster flow
    assert ub[1].mean() > 0.9 * ub[nx // 2].mean()         # and a flat profile
    assert ua[1].mean() < 0.5 * ua[nx // 2].mean()         # (bounce-back: a sheared one)


def test_open_ends_differ_from_the_wrap_only_near_the_ends():
    nx, ny, T = 16, 128, 20
    walls = geometry.channel_walls(nx, ny, barrier=False)
    per, out = _run(nx, ny, walls, T), _run(nx, ny, walls, T, outflow=True)
    d = np.abs(per.f_view("A") - out.f_view("A")).max(axis=(0, 1))
    assert d[:T].max() > 0 and d[-T:].max() > 0            # the ends feel it ...
    assert d[T + 1:ny - T - 1].max() == 0                  # ... and nothing travels faster than one column per step
    z = _run(nx, ny, walls, T, outflow=True, source_col=-1)
    assert np.isfinite(z.f_view("A")).all()
