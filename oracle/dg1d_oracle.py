"""ctypes front end of the CPU oracle (oracle/dg1d_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package fem_b200 never imports this module.

Parity status: pinned against golden vectors produced by the unmodified reference
(tests/golden/*.npz, generator tests/golden/gen_golden.py); see tests/test_oracle_golden.py.
The C sibling's flux family (lf, ecusp, hllc, ausmdv, lfc) is pinned against its flux.c compiled unmodified
(oracle/_ref/libcsib_euler_flux.so, csib_ref_flux below); the Euler Roe flux against PDE::roe_flux of
dg2d/models/euler/pde.h compiled unmodified with stand-in deal.II types (oracle/_ref/libref_dg2d_roe.so, roe_ref_flux).
Unpinned: the scalar inflow/outflow boundary (the C++ twin needs all of deal.II).

Arrays use the reference layout: scalar (nc, nd); Euler (3, nc, nd); float64, C-contiguous.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, 'libdg1d_oracle.so')
MAXND = 8

PDE = {'linear': 0, 'burger': 1, 'euler': 2, 'varadv': 3}
FLUX = {'central': 0, 'upwind': 1, 'roe': 2, 'godunov': 3, 'rusanov': 4,
        'lf': 5, 'ecusp': 6, 'hllc': 7, 'ausmdv': 8, 'lfc': 9}
CSIB_FLUXES = ('lf', 'ecusp', 'hllc', 'ausmdv', 'lfc')   # dg1d/c/euler/src/dg.h:7-11


class OrcOps(C.Structure):
    _fields_ = [('k', C.c_int),
                ('Vf', C.c_double * (MAXND * MAXND)), ('Vg', C.c_double * (MAXND * MAXND)),
                ('bm', C.c_double * MAXND), ('bp', C.c_double * MAXND),
                ('ark', C.c_double * 3), ('brk', C.c_double * 3),
                ('sqrt3', C.c_double), ('isqrt3', C.c_double), ('gamma', C.c_double),
                ('xg', C.c_double * MAXND), ('xmin', C.c_double), ('dx', C.c_double), ('speed_const', C.c_double)]


def build(force=False):
    src = os.path.join(_HERE, 'dg1d_oracle.c')
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(['make', '-C', _HERE, '-s'])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        assert L.orc_sizeof_ops() == C.sizeof(OrcOps)
        dp = C.POINTER(C.c_double)
        op = C.POINTER(OrcOps)
        L.orc_scalar_residual.argtypes = [op, C.c_int, C.c_int, C.c_long, dp, dp]
        L.orc_rk_update.argtypes = [op, C.c_int, C.c_long, dp, dp, dp, C.c_double]
        L.orc_scalar_limiter.argtypes = [op, C.c_long, dp, C.c_double]
        L.orc_scalar_max_speed.argtypes = [C.c_int, C.c_int, C.c_long, dp]
        L.orc_scalar_max_speed.restype = C.c_double
        L.orc_scalar_solve.argtypes = [op, C.c_int, C.c_int, C.c_int, C.c_long, C.c_double, C.c_double, C.c_double,
                                       C.c_double, dp, dp, dp, C.c_long, dp, dp, C.c_long]
        L.orc_scalar_solve.restype = C.c_long
        L.orc_scalar_residual_bc.argtypes = [op, C.c_int, C.c_int, C.c_long, dp, dp, C.c_int, C.c_int, C.c_double, C.c_double]
        L.orc_scalar_limiter_bc.argtypes = [op, C.c_long, dp, C.c_double]
        L.orc_scalar_solve_bc.argtypes = [op, C.c_int, C.c_int, C.c_int, C.c_long, C.c_double, C.c_double, C.c_double, C.c_double,
                                          dp, dp, dp, C.c_long, C.POINTER(C.c_double), C.c_int, C.c_int, dp]
        L.orc_scalar_solve_bc.restype = C.c_long
        L.orc_euler_residual.argtypes = [op, C.c_int, C.c_long, dp, dp]
        L.orc_euler_limiter.argtypes = [op, C.c_int, C.c_long, dp, C.c_double]
        L.orc_euler_max_speed.argtypes = [op, C.c_long, dp]
        L.orc_euler_max_speed.restype = C.c_double
        L.orc_euler_solve.argtypes = [op, C.c_int, C.c_int, C.c_long, C.c_double, C.c_double, C.c_double, C.c_double,
                                      dp, dp, dp, C.c_long, dp, dp, C.c_long]
        L.orc_euler_solve.restype = C.c_long
        L.orc_euler_roe_flux.argtypes = [C.c_double, dp, dp, dp]
        L.orc_euler_rusanov_flux.argtypes = [C.c_double, dp, dp, dp]
        L.orc_euler_csib_flux.argtypes = [C.c_double, C.c_int, dp, dp, dp]
        L.orc_euler_eigmat.argtypes = [C.c_double, dp, dp, dp]
        L.orc_project.argtypes = [op, dp, C.c_long, dp, dp]
        L.orc_error_norm.argtypes = [C.c_int, C.c_int, dp, dp, C.c_long, C.c_double, dp, dp]
        L.orc_error_norm.restype = C.c_double
        L.orc_eval_points.argtypes = [op, C.c_int, C.c_int, dp, C.c_long, dp, dp, dp]
        _lib = L
    return _lib


def _p(a):
    assert a.dtype == np.float64 and a.flags['C_CONTIGUOUS']
    return a.ctypes.data_as(C.POINTER(C.c_double))


# ------------------------------------------------------------------------------------------------
# Basis operators: restatement of dg1d/python/{scalar,euler}/basis.py:4-28 and the operator set-up
# in dgsolver.py:67-99 / euler.py:50-74.  Same expressions in the same order => same bits; the
# goldens store the reference's own operator arrays and tests compare them bit for bit.
# ------------------------------------------------------------------------------------------------
def legendre_table(k, x):
    """P[n], dP[n] for n = 0..k at scalar/array x, by the recurrences of basis.py:4-21."""
    x = np.asarray(x, dtype=np.float64)
    P = [np.ones_like(x), x.copy()]
    dP = [np.zeros_like(x), np.ones_like(x)]
    for n in range(2, k + 1):
        P.append((2.0 * n - 1.0) / n * x * P[n - 1] - (n - 1.0) / n * P[n - 2])
        dP.append(n * P[n - 1] + x * dP[n - 1])
    return P[:k + 1], dP[:k + 1]


def operators(k):
    """Returns dict(xg, wg, Vf, Vg, bm, bp): Vf[q,j] = phi_j(xg_q), Vg[i,q] = phi_i'(xg_q) wg_q."""
    nd = k + 1
    xg, wg = np.polynomial.legendre.leggauss(nd)
    P, dP = legendre_table(k, xg)
    Pm, _ = legendre_table(k, np.float64(-1.0))
    Pp, _ = legendre_table(k, np.float64(+1.0))
    Vf = np.empty((nd, nd)); Vg = np.empty((nd, nd)); bm = np.empty(nd); bp = np.empty(nd)
    for n in range(nd):
        s = np.sqrt(2 * n + 1)
        Vf[:, n] = P[n] * s
        Vg[n, :] = dP[n] * s * wg
        bm[n] = Pm[n] * s
        bp[n] = Pp[n] * s
    return dict(xg=xg, wg=wg, Vf=Vf, Vg=Vg, bm=bm, bp=bp)


ARK = np.array([0.0, 3.0 / 4.0, 1.0 / 3.0])
BRK = 1.0 - ARK            # F6: not the literal 2/3
SQRT3 = np.sqrt(3.0)
ISQRT3 = 1.0 / SQRT3


def varadv_max_speed():
    """varadv.py:15-17, evaluated with numpy exactly as the reference does."""
    x = np.linspace(-1.0, 1.0, 100)
    return float(np.abs(1.0 + (1.0 - x ** 2) ** 5).max())


def make_ops(k, gamma=1.4, Vf=None, Vg=None, bm=None, bp=None, xmin=0.0, dx=0.0):
    o = operators(k)
    Vf = o['Vf'] if Vf is None else Vf
    Vg = o['Vg'] if Vg is None else Vg
    bm = o['bm'] if bm is None else bm
    bp = o['bp'] if bp is None else bp
    nd = k + 1
    s = OrcOps()
    s.k = k
    for q in range(nd):
        for j in range(nd):
            s.Vf[q * nd + j] = Vf[q, j]
            s.Vg[q * nd + j] = Vg[q, j]
    for j in range(nd):
        s.bm[j] = bm[j]
        s.bp[j] = bp[j]
    for r in range(3):
        s.ark[r] = ARK[r]
        s.brk[r] = BRK[r]
    s.sqrt3, s.isqrt3, s.gamma = SQRT3, ISQRT3, gamma
    for j in range(nd):
        s.xg[j] = o['xg'][j]
    s.xmin, s.dx, s.speed_const = float(xmin), float(dx), varadv_max_speed()
    return s


# ------------------------------------------------------------------------------------------------
# thin wrappers
# ------------------------------------------------------------------------------------------------
def scalar_residual(ops, pde, flux, u1):
    res = np.empty_like(u1)
    lib().orc_scalar_residual(C.byref(ops), PDE[pde], FLUX[flux], u1.shape[0], _p(u1), _p(res))
    return res


def rk_update(ops, rk, u0, u1, res, lam):
    """In place on u1 (any shape)."""
    lib().orc_rk_update(C.byref(ops), rk, u1.size, _p(u0), _p(u1), _p(res), float(lam))
    return u1


def scalar_limiter(ops, u1, Mdx2):
    lib().orc_scalar_limiter(C.byref(ops), u1.shape[0], _p(u1), float(Mdx2))
    return u1


def scalar_max_speed(pde, k, u1):
    return lib().orc_scalar_max_speed(PDE[pde], k, u1.shape[0], _p(u1))


def resolve_scalar_flux(pde, num_flux):
    """The reference's flux selection, dgsolver.py:43-59."""
    if pde == 'varadv':
        return 'upwind'          # dgsolver.py:49-51: varadv brings its own numflux
    if pde == 'linear':
        return 'central' if num_flux == 'central' else 'upwind'
    if num_flux == 'central':
        return 'central'
    if num_flux in ('roe', 'upwind'):
        return 'roe'
    return 'godunov'


def scalar_solve(ops, pde, flux, limit, u1, cfl, dx, Tf, Mdx2, max_steps=-1, dt_cap=0):
    """Runs dgsolver.py:255-278 on u1 in place.  cfl is the effective cfl (args.cfl/(2k+1)).
    Returns (it, t, dts)."""
    u0 = np.empty_like(u1); res = np.empty_like(u1)
    t = C.c_double(0.0)
    dts = np.zeros(max(dt_cap, 1))
    it = lib().orc_scalar_solve(C.byref(ops), PDE[pde], FLUX[flux], int(limit), u1.shape[0], cfl, dx, Tf, Mdx2,
                                _p(u0), _p(u1), _p(res), max_steps, C.byref(t), _p(dts), dt_cap)
    return it, t.value, dts[:min(it, dt_cap)]


BC = {'inflow': 3, 'outflow': 1}


def scalar_residual_bc(ops, pde, flux, u1, bcl, bcr, gl=0.0, gr=0.0):
    """Closed ends (the C++ twin's inflow/outflow, dg1d/scalar_legendre/dg.h:367-413; parity unpinned)."""
    res = np.empty_like(u1)
    lib().orc_scalar_residual_bc(C.byref(ops), PDE[pde], FLUX[flux], u1.shape[0], _p(u1), _p(res), BC[bcl], BC[bcr],
                                 float(gl), float(gr))
    return res


def scalar_limiter_bc(ops, u1, Mdx2):
    lib().orc_scalar_limiter_bc(C.byref(ops), u1.shape[0], _p(u1), float(Mdx2))
    return u1


def stage_times(t, dt):
    """Stage times of SSPRK3 as the twin advances them (dg.h:539, :602): t, then a*t + b*(stage_time + dt)."""
    s0 = t
    s1 = ARK[0] * t + BRK[0] * (s0 + dt)
    s2 = ARK[1] * t + BRK[1] * (s1 + dt)
    return s0, s1, s2


def scalar_solve_bc(ops, pde, flux, limit, u1, cfl, dx, Tf, Mdx2, bcl, bcr, boundary, max_steps=-1):
    """boundary(side, t) -> prescribed state ('left'/'right') at time t.  One step per call of the C loop when
    the wave speed depends on the solution (dt is then only known step by step).  Returns (it, t)."""
    u0 = np.empty_like(u1); res = np.empty_like(u1)
    t = C.c_double(0.0)
    it = 0
    const_speed = pde in ('linear', 'varadv')
    ms_const = 1.0 if pde == 'linear' else ops.speed_const
    while t.value < Tf and (max_steps < 0 or it < max_steps):
        ms = ms_const if const_speed else scalar_max_speed(pde, ops.k, u1)
        dt = cfl * dx / ms
        if t.value + dt > Tf:
            dt = Tf - t.value
        tab = np.array([[boundary('left', s), boundary('right', s)] for s in stage_times(t.value, dt)]).reshape(-1)
        it += lib().orc_scalar_solve_bc(C.byref(ops), PDE[pde], FLUX[flux], int(limit), u1.shape[0], cfl, dx, Tf, Mdx2,
                                        _p(u0), _p(u1), _p(res), 1, C.byref(t), BC[bcl], BC[bcr], _p(tab))
    return it, t.value


def euler_residual(ops, flux, u1):
    res = np.empty_like(u1)
    lib().orc_euler_residual(C.byref(ops), FLUX[flux], u1.shape[1], _p(u1), _p(res))
    return res


def euler_limiter(ops, char_lim, u1, Mdx2):
    lib().orc_euler_limiter(C.byref(ops), int(char_lim), u1.shape[1], _p(u1), float(Mdx2))
    return u1


def euler_max_speed(ops, u1):
    return lib().orc_euler_max_speed(C.byref(ops), u1.shape[1], _p(u1))


def euler_solve(ops, flux, char_lim, u1, cfl, dx, Tf, Mdx2, max_steps=-1, dt_cap=0):
    u0 = np.empty_like(u1); res = np.empty_like(u1)
    t = C.c_double(0.0)
    dts = np.zeros(max(dt_cap, 1))
    it = lib().orc_euler_solve(C.byref(ops), FLUX[flux], int(char_lim), u1.shape[1], cfl, dx, Tf, Mdx2,
                               _p(u0), _p(u1), _p(res), max_steps, C.byref(t), _p(dts), dt_cap)
    return it, t.value, dts[:min(it, dt_cap)]


def euler_numflux(flux, ul, ur, gamma=1.4):
    f = np.empty(3)
    ul = np.ascontiguousarray(ul, dtype=np.float64); ur = np.ascontiguousarray(ur, dtype=np.float64)
    if flux in CSIB_FLUXES:
        lib().orc_euler_csib_flux(gamma, FLUX[flux], _p(ul), _p(ur), _p(f))
        return f
    fn = lib().orc_euler_roe_flux if flux == 'roe' else lib().orc_euler_rusanov_flux
    fn(gamma, _p(ul), _p(ur), _p(f))
    return f


_ROE_SO = os.path.join(_HERE, '_ref', 'libref_dg2d_roe.so')
_roe = None


def roe_ref_available():
    return os.path.exists(_ROE_SO)


def roe_ref_flux(ul, ur):
    """The reference's own Roe flux: PDE::roe_flux<2> of dg2d/models/euler/pde.h:209-285 compiled unmodified (against the
    deal.II stand-in types of oracle/dealii_standin), called with normal (1,0) and zero y-momentum; gamma = 1.4."""
    global _roe
    if _roe is None:
        _roe = C.CDLL(_ROE_SO)
        dp = C.POINTER(C.c_double)
        _roe.ref_dg2d_roe_flux.argtypes = [dp, dp, dp]
    a = np.array(ul, dtype=np.float64); b = np.array(ur, dtype=np.float64); f = np.empty(3)
    _roe.ref_dg2d_roe_flux(_p(a), _p(b), _p(f))
    return f


_CSIB_SO = os.path.join(_HERE, '_ref', 'libcsib_euler_flux.so')
_CSIB_NAMES = {'lf': 'LFFlux', 'ecusp': 'ECUSPFlux', 'hllc': 'HLLCFlux', 'ausmdv': 'AUSMDVFlux', 'lfc': 'LFCFlux'}
_csib = None


def csib_available():
    return os.path.exists(_CSIB_SO)


def csib_ref_flux(flux, ul, ur):
    """The C sibling's own flux function (dg1d/c/euler/src/flux.c compiled unmodified into oracle/_ref,
    gamma = 1.4 fixed there).  LFCFlux transforms its arguments in place, so copies are passed."""
    global _csib
    if _csib is None:
        _csib = C.CDLL(_CSIB_SO)
        C.c_int.in_dll(_csib, 'NVAR').value = 3       # dg.h:28, read by the variable-length arrays of flux.c
    a = np.array(ul, dtype=np.float64); b = np.array(ur, dtype=np.float64); f = np.empty(3)
    getattr(_csib, _CSIB_NAMES[flux])(_p(a), _p(b), _p(f))
    return f


def euler_eigmat(u, gamma=1.4):
    R = np.empty((3, 3)); L = np.empty((3, 3))
    u = np.ascontiguousarray(u, dtype=np.float64)
    lib().orc_euler_eigmat(gamma, _p(u), _p(R), _p(L))
    return R, L


def project(ops, wg, vals):
    """vals (nc, Nq) point values at the Gauss points -> modal coefficients (nc, nd)."""
    out = np.empty_like(vals)
    wg = np.ascontiguousarray(wg)
    lib().orc_project(C.byref(ops), _p(wg), vals.shape[0], _p(vals), _p(out))
    return out


def error_norm(k, u1, exact, dx):
    """dgsolver.py:232-252 with ng = k+3; u1 (nc, nd), exact (nc, ng) = the exact solution at the Gauss points."""
    ng = k + 3
    xg, wg = np.polynomial.legendre.leggauss(ng)
    P, _ = legendre_table(k, xg)
    V = np.ascontiguousarray(np.stack([P[n] * np.sqrt(2 * n + 1) for n in range(k + 1)], axis=1))
    return lib().orc_error_norm(k + 1, ng, _p(V), _p(np.ascontiguousarray(wg)), u1.shape[0], float(dx),
                                _p(np.ascontiguousarray(u1)), _p(np.ascontiguousarray(exact)))


def eval_points(ops, Vu, u1):
    """Uniform-point evaluation (dgsolver.py:131-139, euler.py:83-100).  u1 (nc, nd) -> f (nc, nu), or (3, nc, nd) ->
    (r, v, pre) as (3, nc, nu); also the running (min, max) per quantity, shape (nvar, 2)."""
    u1 = np.ascontiguousarray(u1, dtype=np.float64)
    Vu = np.ascontiguousarray(Vu, dtype=np.float64)
    nvar = 1 if u1.ndim == 2 else 3
    nc, nu = u1.shape[-2], Vu.shape[0]
    out = np.empty((nvar, nc, nu)); mm = np.empty((nvar, 2))
    lib().orc_eval_points(C.byref(ops), nvar, nu, _p(Vu), nc, _p(u1), _p(out), _p(mm))
    return (out[0] if nvar == 1 else out), mm


def gauss_points(xmin, dx, nc, xg):
    """x[i,q] as in dgsolver.py:118-119 / euler.py:163-164."""
    i = np.arange(nc)
    xc = xmin + i * dx + 0.5 * dx
    return xc[:, None] + (0.5 * dx * xg)[None, :]
