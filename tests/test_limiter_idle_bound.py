"""The kernels skip the limiter of a cell when a cheap test proves that it cannot change the cell (dg1d_kernels.cuh,
LIM_EPS / eu_char_limiter_idle).  This file restates that test in numpy and checks on the CPU, against the oracle's
limiter (pinned to the reference, euler.py:239-273 / dgsolver.py:206-228), that it never claims "idle" for a cell the
limiter does change -- over ordinary, hypersonic, near-vacuum and badly scaled states and mode-1 amplitudes on both
sides of the 1e-6 threshold -- and that it is not vacuous (it does recognise smooth data)."""
import numpy as np
import pytest

from oracle import dg1d_oracle as orc

LIM_EPS = 1.0e-6
T = LIM_EPS * (1.0 - 2.0 ** -10)


def mid_range(x):
    with np.errstate(all='ignore'):
        return (x >= 2.0 ** -300) & (x < 2.0 ** 300)


def char_idle(gam, uc, du):
    """eu_char_limiter_idle of dg1d_kernels.cuh (a*b+c here where the kernel has fma: the 2^-10 margin covers it)."""
    g1 = gam - 1.0
    g2 = 0.5 * g1
    with np.errstate(all='ignore'):
        d, am = uc[0], np.abs(uc[1])
        a0, a1, a2 = np.abs(du[0]), np.abs(du[1]), np.abs(du[2])
        m2, Ed, d2 = am * am, uc[2] * d, d * d
        Pp = -0.5 * m2 + Ed
        G = (gam * g1) * Pp
        A = (g2 * m2) * a0 + g1 * ((am * d) * a1 + d2 * a2)
        lhs0, rhs0 = G * a0 + A, T * G
        Q = T * d2 - (am * a0 + d * a1)
        rhs1 = G * (Q * Q)
        return (d > 0.0) & (Pp >= 2.0 ** -25 * Ed) & (lhs0 <= rhs0) & (Q > 0.0) & (A * A <= rhs1) & mid_range(rhs0) & mid_range(rhs1)


def states(rng, n, regime):
    """Cell averages (3, n): density, momentum, energy."""
    lu = lambda lo, hi: 10.0 ** rng.uniform(np.log10(lo), np.log10(hi), n)
    sgn = rng.choice([-1.0, 1.0], n)
    if regime == 'ordinary':
        rho, p, mach = lu(0.1, 10.0), lu(0.1, 10.0), rng.uniform(0.0, 3.0, n)
    elif regime == 'hypersonic':      # kinetic energy dominates: p is a small difference of large numbers
        rho, p, mach = lu(0.1, 10.0), lu(1e-9, 1e-2), lu(10.0, 1e5)
    elif regime == 'near_vacuum':
        rho, p, mach = lu(1e-12, 1e-4), lu(1e-14, 1e-4), rng.uniform(0.0, 2.0, n)
    elif regime == 'scaled_up':
        rho, p, mach = lu(1e20, 1e40), lu(1e20, 1e40), rng.uniform(0.0, 3.0, n)
    elif regime == 'scaled_down':
        rho, p, mach = lu(1e-40, 1e-20), lu(1e-40, 1e-20), rng.uniform(0.0, 3.0, n)
    else:                             # at rest: momentum exactly zero
        rho, p, mach = lu(0.1, 10.0), lu(0.1, 10.0), np.zeros(n)
    c = np.sqrt(1.4 * p / rho)
    v = sgn * mach * c
    return np.stack([rho, rho * v, p / 0.4 + 0.5 * rho * v * v])


@pytest.mark.parametrize('regime', ['ordinary', 'hypersonic', 'near_vacuum', 'scaled_up', 'scaled_down', 'rest'])
@pytest.mark.parametrize('k,tvbM', [(1, 0.0), (2, 100.0)])
def test_char_idle_is_conservative(regime, k, tvbM):
    rng = np.random.default_rng(1000 * k + sum(map(ord, regime)))
    n, nd = 200_000, k + 1
    ops = orc.make_ops(k)
    u = np.zeros((3, n, nd))
    u[:, :, 0] = states(rng, n, regime)
    # mode 1: magnitudes from far below to far above the threshold, relative to the size of the characteristic
    # variables (|L du| ~ |du|/rho-ish), every sign pattern, and some exact zeros
    scale = np.abs(u[:, :, 0]).max(axis=0)
    amp = 10.0 ** rng.uniform(-9.5, -3.0, (3, n)) * rng.choice([-1.0, 1.0], (3, n)) * rng.choice([0.0, 1.0, 1.0, 1.0], (3, n))
    u[:, :, 1] = amp * (scale if regime in ('scaled_up', 'scaled_down', 'near_vacuum') else 1.0)
    for m in range(2, nd):
        u[:, :, m] = 0.3 * u[:, :, 1] * rng.standard_normal((3, n))
    # neighbours: smooth (small differences of the averages) on half of the mesh, rough on the other half
    u[:, : n // 2, 0] = np.sort(u[:, : n // 2, 0], axis=1) if regime == 'rest' else u[:, : n // 2, 0]
    dx = 1.0 / n
    idle = char_idle(1.4, u[:, :, 0], u[:, :, 1])
    out = orc.euler_limiter(ops, 1, np.ascontiguousarray(u).copy(), tvbM * dx * dx)
    changed = np.any(out != u, axis=(0, 2))
    assert not np.any(idle & changed), 'a cell the limiter changes was declared idle'
    if regime in ('ordinary', 'rest'):
        assert idle.mean() > 0.05, 'the criterion never fires: vacuous test'


@pytest.mark.parametrize('char_lim', [0, 1])
def test_idle_on_resolved_smooth_data(char_lim):
    """Shu-Osher's initial state at config-4 resolution (dx = 8e-7): everything except the cells at the shock is idle."""
    from fem_b200 import ic as ic_mod
    from fem_b200.basis import Operators
    k, n = 2, 4000
    o = Operators(k)
    ops = orc.make_ops(k)
    dx = 10.0 / 12_500_000
    for x0 in (-4.0 - 2000 * dx, 1.0):          # a window around the shock, a window in the sine region
        x = x0 + (np.arange(n)[:, None] + 0.5) * dx + 0.5 * dx * o.xg[None, :]
        r, m, e = ic_mod.EULER['shuosher'].initial_condition(x.reshape(-1))
        u = np.stack([orc.project(ops, o.wg, np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(n, k + 1))) for v in (r, m, e)])
        if char_lim:
            idle = char_idle(1.4, u[:, :, 0], u[:, :, 1])
        else:
            idle = np.all(np.abs(u[:, :, 1]) <= LIM_EPS, axis=0)
        out = orc.euler_limiter(ops, char_lim, u.copy(), 100.0 * dx * dx)
        changed = np.any(out != u, axis=(0, 2))
        assert not np.any(idle & changed)
        assert idle.sum() >= n - 4


def test_scalar_and_conserved_idle_rule():
    """|du| <= 1e-6 for every component => the cell is untouched (scalar limiter and conserved-variable Euler limiter)."""
    rng = np.random.default_rng(7)
    n, k = 100_000, 2
    ops = orc.make_ops(k)
    u = np.zeros((n, k + 1))
    u[:, 0] = np.cumsum(rng.standard_normal(n)) * 1e-3
    u[:, 1] = 10.0 ** rng.uniform(-9, -4, n) * rng.choice([-1.0, 1.0], n)
    u[:, 2] = 0.5 * u[:, 1]
    # include values exactly at and next to the threshold
    u[:5, 1] = [1e-6, -1e-6, np.nextafter(1e-6, 1), np.nextafter(1e-6, 0), 0.0]
    for M in (0.0, 50.0):
        out = orc.scalar_limiter(ops, u.copy(), M * (1.0 / n) ** 2)
        changed = np.any(out != u, axis=1)
        assert not np.any(changed & (np.abs(u[:, 1]) <= LIM_EPS))
    ue = np.zeros((3, n, k + 1))
    ue[:, :, 0] = states(rng, n, 'ordinary')
    ue[:, :, 1] = 10.0 ** rng.uniform(-9, -4, (3, n)) * rng.choice([-1.0, 1.0], (3, n))
    ue[:, :, 2] = 0.5 * ue[:, :, 1]
    out = orc.euler_limiter(ops, 0, ue.copy(), 0.0)
    changed = np.any(out != ue, axis=(0, 2))
    assert not np.any(changed & np.all(np.abs(ue[:, :, 1]) <= LIM_EPS, axis=0))
