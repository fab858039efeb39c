"""CPU-only checks: the C-ABI library loads and exports every symbol include/dg1d.h declares, argument
validation works without a GPU, and the host-side logic (operators, flux selection, IC plug-ins,
class construction) matches the reference's behaviour recorded in the goldens."""
import ctypes as C
import os
import re
import types

import numpy as np
import pytest

from conftest import ROOT, golden_names, load_golden, rel_linf


@pytest.fixture(scope='module')
def lib():
    import __graft_entry__ as ge
    ge.build()
    from fem_b200 import _lib
    return _lib.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, 'include', 'dg1d.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    declared = set(re.findall(r'\b(dg1d_[a-z0-9_]+)\s*\(', hdr))
    assert len(declared) >= 20
    from fem_b200 import _lib
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.dg1d_version() >= 100


def test_create_validation_without_gpu(lib):
    from fem_b200 import _lib
    h = C.c_void_p()

    def cfg(**kw):
        base = dict(pde=0, flux=1, degree=1, limiter=0, char_lim=0, bc_left=0, bc_right=0, nprob=1, ncell=100,
                    cell_offset=0, ncell_global=100, gamma=1.4)
        base.update(kw)
        return _lib.Config(**base)

    assert lib.dg1d_create(C.byref(cfg()), C.byref(h)) == 0
    assert lib.dg1d_pitch(h) % 16 == 0 and lib.dg1d_pitch(h) >= 100 + 2 * _lib.GHOST
    assert lib.dg1d_state_doubles(h) == 2 * lib.dg1d_pitch(h)
    # compute entry points refuse to run before the operators are set
    assert lib.dg1d_residual(h, 16, 32, None) == -4
    assert b'dg1d_set_operators' in lib.dg1d_last_error()
    lib.dg1d_destroy(h)
    assert lib.dg1d_create(C.byref(cfg(pde=3, flux=1)), C.byref(h)) == 0     # varadv: upwind form only
    lib.dg1d_destroy(h)
    for bad in (dict(pde=7), dict(pde=3, flux=0), dict(degree=5), dict(pde=0, flux=4), dict(pde=1, flux=1), dict(pde=2, flux=1),
                dict(ncell=3), dict(nprob=0), dict(nprob=65536), dict(nprob=100000), dict(bc_left=0, bc_right=1), dict(limiter=1, degree=0),
                dict(pde=2, flux=4, gamma=1.0), dict(cell_offset=50, ncell_global=100)):
        assert lib.dg1d_create(C.byref(cfg(**bad)), C.byref(h)) == -1, bad
        assert len(lib.dg1d_last_error()) > 0


def test_set_operators_validation_without_gpu(lib):
    """The kernels rely on phi_0 = 1, phi_0' = 0 and on SSP-RK3's first stage being u - lam*res: anything else is refused."""
    from fem_b200 import _lib, basis
    h = C.c_void_p()
    cfg = _lib.Config(pde=0, flux=1, degree=2, limiter=0, char_lim=0, bc_left=0, bc_right=0, nprob=1, ncell=100,
                      cell_offset=0, ncell_global=100, gamma=1.4)
    assert lib.dg1d_create(C.byref(cfg), C.byref(h)) == 0
    o = basis.Operators(2)
    dp = lambda a: np.ascontiguousarray(a, dtype=np.float64).ctypes.data_as(C.POINTER(C.c_double))

    def call(Vg=o.Vg, bm=o.bm, ark=basis.ark, brk=basis.brk):
        keep = [np.ascontiguousarray(x, dtype=np.float64) for x in (o.Vf, Vg, bm, o.bp, o.wg, ark, brk)]
        return lib.dg1d_set_operators(h, *[k.ctypes.data_as(C.POINTER(C.c_double)) for k in keep], float(basis.sqrt3), float(basis.isqrt3))

    assert call() == 0
    assert call(ark=np.array([0.25, 0.75, 1.0 / 3.0])) == -1 and b'ark[0]' in lib.dg1d_last_error()
    assert call(brk=np.array([0.5, 0.25, 2.0 / 3.0])) == -1
    bad = np.array(o.Vg, dtype=np.float64).copy(); bad[0, 1] = 1e-300
    assert call(Vg=bad) == -1
    assert call(bm=np.array([1.0 + 1e-15, 1.0, 1.0])) == -1
    lib.dg1d_destroy(h)


def test_argument_validation_of_round2_entry_points_without_gpu(lib):
    """dg1d_eval_points, dg1d_pack_cells / dg1d_unpack_cells and dg1d_stage_part refuse bad arguments before touching the GPU."""
    from fem_b200 import _lib
    h = C.c_void_p()
    cfg = _lib.Config(pde=2, flux=4, degree=2, limiter=1, char_lim=1, bc_left=1, bc_right=1, nprob=1, ncell=100,
                      cell_offset=0, ncell_global=100, gamma=1.4)
    assert lib.dg1d_create(C.byref(cfg), C.byref(h)) == 0
    Vu = (C.c_double * 40)()
    fake = C.c_void_p(256)                     # never dereferenced: every call below fails its argument checks
    for nu in (0, 9, -1):
        assert lib.dg1d_eval_points(h, fake, Vu, nu, fake, None, None) == -1 and b'nu' in lib.dg1d_last_error()
    assert lib.dg1d_eval_points(h, None, Vu, 3, fake, None, None) == -1
    for c0, n in ((-1, 10), (0, 0), (95, 6), (100, 1)):
        assert lib.dg1d_pack_cells(h, fake, fake, c0, n, None) == -1 and b'cell range' in lib.dg1d_last_error()
        assert lib.dg1d_unpack_cells(h, fake, fake, c0, n, None) == -1
    # the operators are not set: the stage entry points say so (and so never launch)
    assert lib.dg1d_stage_part(h, 1 | (1 << 4), 0, fake, None, fake, fake, None, 0.0, None, None) == -4
    lib.dg1d_destroy(h)


def test_euler_flux_family_and_peer_transport_validation(lib):
    """The C sibling's flux ids are Euler-only; the peer-transport entry points reject null handles without a GPU."""
    from fem_b200 import _lib
    h = C.c_void_p()
    base = dict(pde=2, flux=4, degree=2, limiter=1, char_lim=1, bc_left=1, bc_right=1, nprob=1, ncell=100,
                cell_offset=0, ncell_global=100, gamma=1.4)
    for name in _lib.EULER_FLUXES:
        cfg = _lib.Config(**dict(base, flux=_lib.FLUX[name]))
        assert lib.dg1d_create(C.byref(cfg), C.byref(h)) == 0, name
        lib.dg1d_destroy(h)
    for pde in (0, 1, 3):
        for f in (5, 6, 7, 8, 9):
            cfg = _lib.Config(**dict(base, pde=pde, flux=f, limiter=0, char_lim=0, bc_left=0, bc_right=0))
            assert lib.dg1d_create(C.byref(cfg), C.byref(h)) == -1
    assert lib.dg1d_create(C.byref(_lib.Config(**dict(base, flux=10))), C.byref(h)) == -1
    buf = C.create_string_buffer(64)
    n = C.c_int64(-1)
    assert lib.dg1d_dist_p2p_open(None, None, buf) == -1
    assert lib.dg1d_dist_p2p_connect(None, buf) == -1
    assert lib.dg1d_dist_p2p_error(None, C.byref(n)) == -1
    assert lib.dg1d_dist_transport(None) == 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from fem_b200 import _lib
    from fem_b200.device import Dg1dDevice
    with pytest.raises(_lib.Dg1dError):
        Dg1dDevice('linear', 'upwind', 1, 100)


@pytest.mark.parametrize('name', golden_names())
def test_product_operators_bitwise(name):
    from fem_b200 import basis
    g = load_golden(name)
    o = basis.Operators(g['degree'])
    for key in ('xg', 'wg', 'Vf', 'Vg', 'Vu', 'bm', 'bp'):
        assert np.array_equal(getattr(o, key), g[key]), key
    assert np.array_equal(basis.ark, g['ark']) and np.array_equal(basis.brk, g['brk'])
    assert basis.brk[2].hex() == '0x1.5555555555556p-1'
    if g['kind'] == 'euler':
        assert basis.sqrt3 == g['sqrt3'] and basis.isqrt3 == g['isqrt3']


def test_basis_api():
    from fem_b200 import basis
    x = np.array([-0.3, 0.2, 0.9])
    assert np.allclose(basis.Legendre(3, x), 0.5 * (5 * x ** 3 - 3 * x))
    assert np.allclose(basis.dLegendre(3, x), 0.5 * (15 * x ** 2 - 3))
    assert np.allclose(basis.shape_value(2, x), np.sqrt(5) * 0.5 * (3 * x ** 2 - 1))
    V = basis.Vandermonde(2, x)
    assert V.shape == (3, 3) and np.allclose(V[:, 1], np.sqrt(3) * x)
    # mass matrix of the basis is 2*I on [-1,1] (so dx*I on a cell)
    o = basis.Operators(4)
    M = (o.Vf * o.wg[:, None]).T @ o.Vf
    assert np.allclose(M, 2 * np.eye(5), atol=1e-14)


def test_flux_resolution_follows_reference():
    from fem_b200.device import resolve_scalar_flux as r
    assert r('linear', 'central') == 'central'
    for f in ('upwind', 'roe', 'godunov'):
        assert r('linear', f) == 'upwind'                 # dgsolver.py:45-48
    assert r('burger', 'central') == 'central'
    assert r('burger', 'roe') == 'roe' and r('burger', 'upwind') == 'roe'   # dgsolver.py:56-57
    assert r('burger', 'godunov') == 'godunov'


def _ns(**kw):
    base = dict(pde='linear', ncell=50, nrefine=0, degree=1, cfl=0.9, Tf=1.0, plot_freq=0, ic='sin2pi', limit='no',
                tvbM=0.0, compute_error='no', num_flux='upwind')
    base.update(kw)
    return types.SimpleNamespace(**base)


def test_dg_constructor_mirrors_reference():
    from fem_b200.dgsolver import DG
    d = DG(_ns(degree=2, ncell=80, tvbM=10.0, ic='mult', pde='burger', num_flux='upwind'))
    assert d.k == 2 and d.nd == 3 and d.nc == 80
    assert d.cfl == 0.9 / 5 and d.xmin == -1.0 and d.xmax == 1.0
    assert d.dx == 2.0 / 80 and d.Mdx2 == 10.0 * d.dx ** 2
    assert d.num_flux == 'roe'
    with pytest.raises(SystemExit):
        DG(_ns(ic='nope'))
    with pytest.raises(SystemExit):
        DG(_ns(pde='nope'))


def test_euler_constructor_and_cli_flags():
    from fem_b200.euler import Euler, make_parser
    a = make_parser().parse_args([])
    assert (a.ncell, a.degree, a.cfl, a.Tf, a.ic, a.char_lim, a.tvbM) == (100, 1, 0.9, 0.0, 'sod', 0, 0.0)
    e = Euler(a)
    assert e.Tf == 0.2 and e.cfl == 0.3 and e.dx == 0.01 and e.num_flux == 'rusanov'   # euler.py:177-178: Tf from the IC
    e = Euler(make_parser().parse_args(['-ic', 'shuosher', '-Tf', '0.5', '-degree', '2', '-tvbM', '100']))
    assert e.Tf == 0.5 and e.xmin == -5.0 and abs(e.Mdx2 - 100 * 0.1 ** 2) < 1e-15
    for f in ('roe', 'lf', 'ecusp', 'hllc', 'ausmdv', 'lfc'):       # extension: dg1d/c/euler/src/dg.h:7-11
        assert Euler(make_parser().parse_args(['-num_flux', f])).num_flux == f
    with pytest.raises(SystemExit):
        make_parser().parse_args(['-num_flux', 'godunov'])
    from fem_b200.dg import make_parser as mp
    a = mp().parse_args([])
    assert (a.pde, a.ncell, a.degree, a.cfl, a.Tf, a.ic, a.limit, a.num_flux) == ('linear', 50, 1, 0.9, 1.0, 'sin2pi', 'no',
                                                                                  'upwind')


@pytest.mark.parametrize('name', golden_names())
def test_ic_plugins_reproduce_reference_projection(name):
    """IC plug-ins evaluated in one vectorised call + the oracle's projection == the reference's u_init."""
    from fem_b200 import ic as ic_mod
    from oracle import dg1d_oracle as orc
    g = load_golden(name)
    ops = orc.make_ops(g['degree'])
    x = ic_mod.gauss_points(g['xmin'], g['dx'], g['ncell'], g['xg'])
    nc, nq = x.shape
    if g['kind'] == 'scalar':
        vals = ic_mod.SCALAR[g['ic']].initial_condition(x.reshape(-1)).reshape(nc, nq)
        assert rel_linf(orc.project(ops, g['wg'], np.ascontiguousarray(vals)), g['u_init']) <= 1e-14
    else:
        if 'riemann_vl' in g:
            plug = ic_mod.riemann_plugin(g['riemann_vl'], g['riemann_vr'], float(g['riemann_xs']))
        else:
            plug = ic_mod.EULER[g['ic']]
        for v, vals in enumerate(plug.initial_condition(x.reshape(-1))):
            u = orc.project(ops, g['wg'], np.ascontiguousarray(np.asarray(vals, dtype=np.float64).reshape(nc, nq)))
            assert rel_linf(u, g['u_init'][v]) <= 1e-14


def test_dirichlet_option_host_logic():
    """-bc dirichlet (the C++ twin's exercise, dg1d/scalar_legendre/README.md:91-137): constructor checks, the default
    boundary data, and the stage times the host evaluates them at (dg.h:539, :602), without a GPU."""
    from fem_b200.basis import ark, brk
    from fem_b200.device import Dg1dDevice
    from fem_b200.dg import make_parser
    from fem_b200.dgsolver import DG
    from oracle import dg1d_oracle as orc
    assert make_parser().parse_args([]).bc == 'periodic'
    assert DG(_ns()).bc == 'periodic'                        # a reference-style namespace without the attribute
    d = DG(_ns(bc='dirichlet'))
    t = 0.3
    assert d.boundary_condition(0.0, t) == float(1 + np.sin(2 * np.pi * np.array([0.0 - t]))[0])   # ic(x - t), unit speed
    b = DG(_ns(bc='dirichlet', pde='burger', num_flux='roe'))
    assert b.boundary_condition(0.0, t) == b.boundary_condition(0.0, 0.0)                          # held at its initial value
    with pytest.raises(SystemExit):
        DG(_ns(bc='neumann'))
    for t0, dt in ((0.0, 0.01), (0.37, 1.0 / 3.0), (1.25, 7e-4)):
        s = Dg1dDevice.stage_times(t0, dt, ark, brk)
        assert s == orc.stage_times(t0, dt)
        assert s[0] == t0 and s[1] == t0 + dt and abs(s[2] - (t0 + 0.5 * dt)) <= 2e-16 * max(1.0, t0)
