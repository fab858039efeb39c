"""Pins the CPU oracle (oracle/) against golden vectors produced by the unmodified reference.

CPU-only.  Every stage piece (residual, RK update, limiter, max_speed/dt) is compared on the
reference's own stage inputs, and the full solve is compared at final time.  Tolerances are the
ones north_star states: 1e-12 relative L-inf per stage, 1e-10 at final time; the observed
differences are recorded in DESIGN.md (bit-identical for most k <= 2 cases).
"""
import numpy as np
import pytest

from conftest import golden_names, load_golden, rel_linf
from oracle import dg1d_oracle as orc

STAGE_TOL = 1e-12
FINAL_TOL = 1e-10


@pytest.mark.parametrize('name', golden_names())
def test_operators_bitwise(name):
    g = load_golden(name)
    o = orc.operators(g['degree'])
    for key in ('xg', 'wg', 'Vf', 'bm', 'bp'):
        assert np.array_equal(o[key], g[key]), key
    # the reference stores Vg[i,j] = phi_i'(xg_j) wg_j in both solvers
    assert np.array_equal(o['Vg'], g['Vg'])
    assert np.array_equal(orc.ARK, g['ark']) and np.array_equal(orc.BRK, g['brk'])
    assert orc.BRK[2].hex() == '0x1.5555555555556p-1'
    if g['kind'] == 'euler':
        assert orc.SQRT3 == g['sqrt3'] and orc.ISQRT3 == g['isqrt3']


@pytest.mark.parametrize('name', golden_names('scalar'))
def test_scalar_stages(name):
    g = load_golden(name)
    k = g['degree']
    ops = orc.make_ops(k, xmin=g['xmin'], dx=g['dx'])
    flux = orc.resolve_scalar_flux(g['pde'], g['num_flux'])
    limit = g['limit'] == 'yes'
    worst = 0.0
    for s in range(len(g['stage_id'])):
        u_in = np.ascontiguousarray(g['stage_in'][s])
        res = orc.scalar_residual(ops, g['pde'], flux, u_in)
        worst = max(worst, rel_linf(res, g['res'][s]))
        assert rel_linf(res, g['res'][s]) <= STAGE_TOL
        if limit:
            # limiter alone, on the reference's own pre-limiter state
            pre = np.ascontiguousarray(g['pre_lim'][s]).copy()
            post = orc.scalar_limiter(ops, pre, g['Mdx2'])
            assert rel_linf(post, g['post_lim'][s]) <= STAGE_TOL
    print(name, 'worst residual rel err', worst)


@pytest.mark.parametrize('name', golden_names('scalar'))
def test_scalar_final(name):
    g = load_golden(name)
    k = g['degree']
    ops = orc.make_ops(k, xmin=g['xmin'], dx=g['dx'])
    flux = orc.resolve_scalar_flux(g['pde'], g['num_flux'])
    u1 = np.ascontiguousarray(g['u_init']).copy()
    it, t, _ = orc.scalar_solve(ops, g['pde'], flux, g['limit'] == 'yes', u1, g['cfl_eff'], g['dx'], g['Tf'],
                                g['Mdx2'])
    assert 3 * it == g['nstages']
    err = rel_linf(u1, g['u_final'])
    print(name, 'final rel err', err, 'bitwise' if np.array_equal(u1, g['u_final']) else '')
    assert err <= FINAL_TOL


@pytest.mark.parametrize('name', golden_names('scalar'))
def test_scalar_projection(name):
    g = load_golden(name)
    ops = orc.make_ops(g['degree'])
    ic = SCALAR_IC[g['ic']]
    x = orc.gauss_points(g['xmin'], g['dx'], g['ncell'], g['xg'])
    u = orc.project(ops, g['wg'], np.ascontiguousarray(ic(x)))
    assert rel_linf(u, g['u_init']) <= 1e-14


@pytest.mark.parametrize('name', golden_names('scalar'))
def test_scalar_error_norm(name):
    g = load_golden(name)
    k = g['degree']
    xg, _ = np.polynomial.legendre.leggauss(k + 3)
    x = orc.gauss_points(g['xmin'], g['dx'], g['ncell'], xg)
    err = orc.error_norm(k, np.ascontiguousarray(g['u_final']), np.ascontiguousarray(SCALAR_IC[g['ic']](x)), g['dx'])
    assert abs(err - g['err']) <= 1e-12 * max(g['err'], 1e-300)


def _hat(x):
    return np.where((x < 0.25) | (x > 0.75), 0.0, 1.0)


def _mult(x):
    f = np.zeros_like(x)
    m = (x > -0.8) & (x < -0.6); f[m] = np.exp(-np.log(2.0) * (x[m] + 0.7) ** 2 / 0.0009)
    m = (x > -0.4) & (x < -0.2); f[m] = 1.0
    m = (x > 0.0) & (x < 0.2); f[m] = 1.0 - np.abs(10.0 * (x[m] - 0.1))
    m = (x > 0.4) & (x < 0.6); f[m] = np.sqrt(1.0 - 100.0 * (x[m] - 0.5) ** 2)
    return f


SCALAR_IC = {
    'sin2pi': lambda x: 1 + np.sin(2 * np.pi * x),
    'sin4pi': lambda x: np.sin(4 * np.pi * x),
    'gauss': lambda x: np.exp(-100 * (x - 0.5) ** 2),
    'hat': _hat,
    'mult': _mult,
}


@pytest.mark.parametrize('name', golden_names('euler'))
def test_euler_stages(name):
    g = load_golden(name)
    k = g['degree']
    ops = orc.make_ops(k)
    worst = [0.0, 0.0, 0.0]
    for s in range(len(g['stage_id'])):
        sid = int(g['stage_id'][s]); step, rk = divmod(sid, 3)
        u_in = np.ascontiguousarray(g['stage_in'][s])
        if not np.all(np.isfinite(u_in)):
            continue
        res = orc.euler_residual(ops, g.get('num_flux', 'rusanov'), u_in)
        e = rel_linf(res, g['res'][s]); worst[0] = max(worst[0], e)
        assert e <= STAGE_TOL, (name, s, 'residual', e)
        # RK update on the reference's own residual
        if rk == 0:
            u0 = u_in
        else:
            j = list(g['stage_id']).index(sid - rk)
            u0 = np.ascontiguousarray(g['stage_in'][j])
        u1 = u_in.copy()
        orc.rk_update(ops, rk, u0, u1, np.ascontiguousarray(g['res'][s]), g['lam'][step])
        e = rel_linf(u1, g['pre_lim'][s]); worst[1] = max(worst[1], e)
        assert np.array_equal(u1, g['pre_lim'][s]), (name, s, 'update not bitwise', e)
        pre = np.ascontiguousarray(g['pre_lim'][s]).copy()
        post = orc.euler_limiter(ops, g['char_lim'], pre, g['Mdx2'])
        e = rel_linf(post, g['post_lim'][s]); worst[2] = max(worst[2], e)
        assert e <= STAGE_TOL, (name, s, 'limiter', e)
    print(name, 'worst rel err: residual %.2e update %.2e limiter %.2e' % tuple(worst))


@pytest.mark.parametrize('name', golden_names('euler'))
def test_euler_dt_sequence_and_final(name):
    g = load_golden(name)
    k = g['degree']
    ops = orc.make_ops(k)
    u1 = np.ascontiguousarray(g['u_init']).copy()
    assert orc.euler_max_speed(ops, u1) > 0
    ncap = len(g['dt'])
    it, t, dts = orc.euler_solve(ops, g.get('num_flux', 'rusanov'), g['char_lim'], u1, g['cfl_eff'], g['dx'], g['Tf'], g['Mdx2'],
                                 dt_cap=ncap)
    if not np.isfinite(g['t']):
        # the reference blew up (negative pressure -> NaN dt ends its loop); the oracle must too
        assert not np.isfinite(t)
        return
    assert it == g['it']
    assert rel_linf(dts, g['dt']) <= FINAL_TOL
    assert abs(t - g['t']) <= 1e-14
    err = rel_linf(u1, g['u_final'])
    print(name, 'final rel err', err, 'bitwise' if np.array_equal(u1, g['u_final']) else '')
    assert err <= FINAL_TOL
    # conservation: the oracle's totals equal the reference's; where the reference's end cells are still
    # undisturbed (copy BC), mass is constant and momentum grows by (p_l - p_r) t to round-off
    dx = g['dx']
    for v in range(3):
        ref = g['u_final'][v, :, 0].sum()
        assert abs(u1[v, :, 0].sum() - ref) <= 1e-11 * max(1.0, abs(ref))
    gi, gf = g['u_init'], g['u_final']
    if g['ic'] == 'sod' and 'riemann_vl' not in g and g.get('num_flux') != 'hllc' and np.array_equal(gi[:, [0, -1], 0], gf[:, [0, -1], 0]):
        assert abs(u1[0, :, 0].sum() * dx - gi[0, :, 0].sum() * dx) < 1e-13
        assert abs(u1[1, :, 0].sum() * dx - (1.0 - 0.1) * t) < 1e-13


@pytest.mark.parametrize('name', [n for n in golden_names('euler') if 'ens' in n])
def test_euler_projection_riemann(name):
    g = load_golden(name)
    ops = orc.make_ops(g['degree'])
    x = orc.gauss_points(g['xmin'], g['dx'], g['ncell'], g['xg'])
    vl, vr, xs = g['riemann_vl'], g['riemann_vr'], g['riemann_xs']
    gam = 1.4
    enel = vl[2] / (gam - 1.0) + 0.5 * vl[0] * vl[1] ** 2
    ener = vr[2] / (gam - 1.0) + 0.5 * vr[0] * vr[1] ** 2
    rho = (x <= xs) * vl[0] + (x > xs) * vr[0]
    mom = (x <= xs) * vl[0] * vl[1] + (x > xs) * vr[0] * vr[1]
    ene = (x <= xs) * enel + (x > xs) * ener
    for v, vals in enumerate((rho, mom, ene)):
        u = orc.project(ops, g['wg'], np.ascontiguousarray(vals))
        assert rel_linf(u, g['u_init'][v]) <= 1e-14


# ---- the C sibling's flux family: the restatement against flux.c itself ---------------------------
@pytest.mark.parametrize('flux', orc.CSIB_FLUXES)
def test_csib_flux_vectors(flux):
    """Vectors recorded from dg1d/c/euler/src/flux.c compiled unmodified (tests/golden/gen_golden.py)."""
    g = load_golden('csibvec_flux_pairs')
    for ul, ur, want in zip(g['ul'], g['ur'], g[flux]):
        got = orc.euler_numflux(flux, ul, ur)
        assert np.array_equal(got, want, equal_nan=True), (flux, ul, ur, got, want)


@pytest.mark.skipif(not orc.csib_available(), reason='oracle/_ref not built (make -C oracle, needs /root/reference)')
@pytest.mark.parametrize('flux', orc.CSIB_FLUXES)
def test_csib_flux_against_compiled_reference(flux):
    """Fresh random pairs against the compiled reference flux itself, bit for bit."""
    rng = np.random.default_rng(orc.CSIB_FLUXES.index(flux) + 5)
    for i in range(3000):
        d = rng.uniform(0.05, 4.0, 2); p = rng.uniform(0.02, 4.0, 2)
        v = rng.uniform(-3.0, 3.0, 2) * np.sqrt(1.4 * p / d) * (i % 4 != 0)
        ul = np.array([d[0], d[0] * v[0], p[0] / 0.4 + 0.5 * d[0] * v[0] ** 2])
        ur = np.array([d[1], d[1] * v[1], p[1] / 0.4 + 0.5 * d[1] * v[1] ** 2]) if i % 3 else ul.copy()
        assert np.array_equal(orc.euler_numflux(flux, ul, ur), orc.csib_ref_flux(flux, ul, ur), equal_nan=True), (flux, ul, ur)


def test_roe_flux_vectors():
    """The Euler Roe flux against vectors recorded from the reference's own PDE::roe_flux (dg2d/models/euler/pde.h:209-285,
    compiled unmodified against deal.II stand-in types: oracle/ref_dg2d_roe.cc), one space dimension."""
    g = load_golden('csibvec_flux_pairs')
    for ul, ur, want in zip(g['ul'], g['ur'], g['roe']):
        assert np.array_equal(orc.euler_numflux('roe', ul, ur), want, equal_nan=True), (ul, ur)


@pytest.mark.skipif(not orc.roe_ref_available(), reason='oracle/_ref not built (make -C oracle, needs /root/reference)')
def test_roe_flux_against_compiled_reference():
    rng = np.random.default_rng(11)
    for i in range(5000):
        d = rng.uniform(0.05, 4.0, 2); p = rng.uniform(0.02, 4.0, 2)
        v = rng.uniform(-3.0, 3.0, 2) * np.sqrt(1.4 * p / d) * (i % 4 != 0)
        ul = np.array([d[0], d[0] * v[0], p[0] / 0.4 + 0.5 * d[0] * v[0] ** 2])
        ur = np.array([d[1], d[1] * v[1], p[1] / 0.4 + 0.5 * d[1] * v[1] ** 2]) if i % 3 else ul * (1 + 1e-4 * rng.standard_normal(3))
        assert np.array_equal(orc.euler_numflux('roe', ul, ur), orc.roe_ref_flux(ul, ur), equal_nan=True), (ul, ur)


@pytest.mark.parametrize('flux', ['lf', 'ecusp', 'ausmdv', 'lfc'])
def test_csib_flux_consistency(flux):
    """F(u,u) = f(u) to round-off for the consistent members of the family (the sibling's HLLC is not: DESIGN.md)."""
    rng = np.random.default_rng(3)
    for _ in range(200):
        d, p = rng.uniform(0.1, 3.0, 2); v = rng.uniform(-2.0, 2.0)
        u = np.array([d, d * v, p / 0.4 + 0.5 * d * v * v])
        exact = np.array([u[1], p + u[1] * v, (u[2] + p) * v])
        assert np.max(np.abs(orc.euler_numflux(flux, u, u) - exact)) <= 1e-13 * np.max(np.abs(exact) + 1.0)


# ---- closed ends for the scalar PDEs: the C++ twin's inflow / outflow (parity unpinned) ---------------
def _closed_run(k, nc, pde='linear', flux='upwind', limit=False, Tf=1.0):
    dx = 1.0 / nc
    ops = orc.make_ops(k, xmin=0.0, dx=dx); o = orc.operators(k)
    ic = SCALAR_IC['sin2pi']
    u = orc.project(ops, o['wg'], np.ascontiguousarray(ic(orc.gauss_points(0.0, dx, nc, o['xg']))))
    it, t = orc.scalar_solve_bc(ops, pde, flux, limit, u, 0.9 / (2 * k + 1), dx, Tf, 0.0, 'inflow', 'outflow',
                                lambda side, s: float(ic(np.array([0.0 - s]))[0]))
    xe = orc.gauss_points(0.0, dx, nc, np.polynomial.legendre.leggauss(k + 3)[0])
    return it, t, orc.error_norm(k, u, ic(xe - t), dx), u


@pytest.mark.parametrize('k,rate', [(1, 2.0), (2, 3.0)])
def test_inflow_outflow_converges_to_the_exact_solution(k, rate):
    """The README exercise of the twin (dg1d/scalar_legendre/README.md:91-137): u_t + u_x = 0 on (0,1), inflow data from the
    exact solution, outflow on the right: order k+1, and to a non-integer final time (no periodicity to lean on)."""
    errs = [_closed_run(k, nc, Tf=0.73)[2] for nc in (20, 40, 80)]
    rates = [np.log2(errs[i] / errs[i + 1]) for i in range(2)]
    assert all(abs(r - rate) < 0.15 for r in rates), (errs, rates)


def test_closed_mesh_matches_the_periodic_one_away_from_the_ends():
    """Same volume term, traces, fluxes and scatter order as the pinned periodic residual: on a state that IS periodic, and
    with the inflow value equal to the periodic neighbour's trace, the interior must agree bit for bit with the periodic
    residual; cell 0 agrees to round-off (upwind flux: the left face only sees the left state, here a re-rounded trace)
    and the last cell differs only by the periodic loop's own scatter-order exception (dgsolver.py:191)."""
    k, nc = 2, 64
    dx = 1.0 / nc
    ops = orc.make_ops(k, xmin=0.0, dx=dx); o = orc.operators(k)
    rng = np.random.default_rng(7)
    u = rng.standard_normal((nc, k + 1))
    per = orc.scalar_residual(ops, 'linear', 'upwind', u)
    gl = float(u[-1] @ o['bp'])          # the periodic left state of face 0 (numpy's dot order: last-ulp agreement only)
    clo = orc.scalar_residual_bc(ops, 'linear', 'upwind', u, 'inflow', 'outflow', gl=gl)
    assert np.array_equal(clo[1:-1], per[1:-1])
    assert rel_linf(clo[0], per[0]) < 1e-13 and rel_linf(clo[-1], per[-1]) < 1e-13


def test_closed_mesh_limiter_leaves_the_end_cells():
    k, nc = 2, 40
    ops = orc.make_ops(k, xmin=0.0, dx=1.0 / nc)
    rng = np.random.default_rng(8)
    u = rng.standard_normal((nc, k + 1))
    v = orc.scalar_limiter_bc(ops, u.copy(), 0.0)
    assert np.array_equal(v[0], u[0]) and np.array_equal(v[-1], u[-1])
    w = orc.scalar_limiter(ops, u.copy(), 0.0)
    assert np.array_equal(v[1:-1], w[1:-1])                  # interior cells: the pinned periodic limiter


def test_closed_mesh_burgers_outflow_both_fluxes_consistent():
    """numflux(u, u) is the physical flux for every scalar flux, so 'outflow' through the copy face is the twin's rule."""
    k, nc = 1, 32
    ops = orc.make_ops(k, xmin=0.0, dx=1.0 / nc)
    rng = np.random.default_rng(9)
    u = 1.5 + 0.3 * rng.standard_normal((nc, k + 1))
    a = orc.scalar_residual_bc(ops, 'burger', 'roe', u, 'inflow', 'outflow', gl=1.2)
    b = orc.scalar_residual_bc(ops, 'burger', 'godunov', u, 'inflow', 'outflow', gl=1.2)
    assert np.array_equal(a[[0, -1]], b[[0, -1]]) or rel_linf(a[[0, -1]], b[[0, -1]]) < 1e-15
