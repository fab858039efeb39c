"""Uniform-point evaluation: the values the reference hands to its plotting calls and its axis limits.
Goldens tests/golden/evalpts_*.npz are recorded from the unmodified reference run with plotting ON under a recording
matplotlib stub (tests/golden/gen_eval_golden.py).  CPU: the oracle against the goldens.  GPU: dg1d_eval_points through the
C ABI against the oracle (bitwise) and the goldens (bitwise for degree <= 2, 1e-12 above: numpy's gemv order)."""
import numpy as np
import pytest

from conftest import golden_names, load_golden, rel_linf


def _want(g):
    if g['kind'] == 'scalar':
        return [('u_init', g['f_init'][None], g['axis_init'][None, 2:], 0.1), ('u_final', g['f_final'][None], g['axis_final'][None, 2:], 0.1)]
    return [('u_final', np.stack([g['r_final'], g['v_final'], g['p_final']]), g['ylim_final'], 0.05)]


@pytest.mark.parametrize('name', golden_names('evalpts'))
def test_oracle_eval_points_against_reference(name):
    from oracle import dg1d_oracle as orc
    g = load_golden(name)
    k = int(g['degree'])
    ops = orc.make_ops(k)
    for key, vals, lims, pad in _want(g):
        out, mm = orc.eval_points(ops, g['Vu'], g[key])
        out = out[None] if out.ndim == 2 else out
        if k <= 2:
            assert np.array_equal(out, vals), (name, key)
        assert rel_linf(out, vals) <= 1e-12
        # the reference's axis limits are umin - pad, umax + pad of ITS values
        assert np.allclose(mm[:, 0] - pad, lims[:, 0], rtol=0, atol=1e-12) and np.allclose(mm[:, 1] + pad, lims[:, 1], rtol=0, atol=1e-12)
        if k <= 2:
            assert np.array_equal(mm[:, 0] - pad, lims[:, 0]) and np.array_equal(mm[:, 1] + pad, lims[:, 1])


@pytest.mark.gpu
@pytest.mark.parametrize('name', golden_names('evalpts'))
def test_device_eval_points(name):
    from oracle import dg1d_oracle as orc
    from fem_b200.device import Dg1dDevice
    g = load_golden(name)
    k, nc = int(g['degree']), int(g['ncell'])
    ops = orc.make_ops(k)
    if g['kind'] == 'scalar':
        dev = Dg1dDevice('linear', 'upwind', k, nc, bc=('periodic', 'periodic'))
    else:
        dev = Dg1dDevice('euler', 'rusanov', k, nc, limiter=(k > 0), bc=('copy', 'copy'))
    assert np.array_equal(dev.ops.Vu, g['Vu']) and np.array_equal(dev.ops.xu, g['xu'])
    for key, vals, lims, pad in _want(g):
        u = np.ascontiguousarray(g[key])
        dev.pack(u.reshape(dev.nvar, 1, nc, k + 1))
        out, mm = dev.eval_points()
        out = out.cpu().numpy().reshape(dev.nvar, nc, -1); mm = mm.cpu().numpy().reshape(dev.nvar, 2)
        oo, om = orc.eval_points(ops, g['Vu'], u)
        oo = oo[None] if oo.ndim == 2 else oo
        assert np.array_equal(out, oo) and np.array_equal(mm, om), 'device differs from the oracle'
        assert rel_linf(out, vals) <= 1e-12
        if k <= 2:
            assert np.array_equal(out, vals)
            assert np.array_equal(mm[:, 0] - pad, lims[:, 0]) and np.array_equal(mm[:, 1] + pad, lims[:, 1])


@pytest.mark.gpu
def test_eval_points_ragged_ensemble_and_nan():
    """Many blocks with a ragged last one, several problems, no min/max requested, and NaN propagation into the limits."""
    import torch
    from oracle import dg1d_oracle as orc
    from fem_b200.device import Dg1dDevice
    rng = np.random.default_rng(3)
    k, nc, nprob = 2, 1000 + 37, 3
    dev = Dg1dDevice('euler', 'rusanov', k, nc, limiter=True, bc=('copy', 'copy'), nprob=nprob)
    u = rng.uniform(0.5, 2.0, (3, nprob, nc, k + 1))
    u[:, :, :, 1:] *= 0.05
    u[1] -= 1.0          # negative velocities too
    u[0, 1, 500, 0] = np.nan
    dev.pack(u)
    out, mm = dev.eval_points()
    out, mm = out.cpu().numpy(), mm.cpu().numpy()
    ops = orc.make_ops(k)
    for p in range(nprob):
        oo, om = orc.eval_points(ops, dev.ops.Vu, np.ascontiguousarray(u[:, p]))
        assert np.array_equal(out[:, p], oo, equal_nan=True)
        assert np.array_equal(mm[:, p], om, equal_nan=True)
    assert np.isnan(mm[:, 1]).all() and np.isfinite(mm[:, 0]).all() and np.isfinite(mm[:, 2]).all()
    out2, none = dev.eval_points(with_minmax=False)
    assert none is None and np.array_equal(out2.cpu().numpy(), out, equal_nan=True)
    # a different set of points: the cell's two ends only
    from fem_b200.basis import Vandermonde
    V2 = Vandermonde(k, np.array([-1.0, 1.0]))
    o3, _ = dev.eval_points(Vu=V2)
    oo, _ = orc.eval_points(ops, V2, np.ascontiguousarray(u[:, 0]))
    assert np.array_equal(o3.cpu().numpy()[:, 0], oo)


@pytest.mark.gpu
def test_class_level_eval_points():
    """DG.eval_points() / Euler.eval_points() after run() against the reference's final plot data."""
    import types
    from fem_b200.dgsolver import DG
    from fem_b200.euler import Euler, make_parser
    g = load_golden('evalpts_bur_k2_nc100_M10')
    dg = DG(types.SimpleNamespace(pde='burger', ncell=100, nrefine=0, degree=2, cfl=0.9, Tf=0.3, plot_freq=0, ic='sin2pi',
                                  limit='yes', tvbM=10.0, compute_error='no', num_flux='roe'))
    dg.run()
    x, f, (umin, umax) = dg.eval_points()
    assert np.array_equal(x, g['x']) and rel_linf(f, g['f_final']) <= 1e-10
    assert abs(umin - 0.1 - g['axis_final'][2]) <= 1e-10 and abs(umax + 0.1 - g['axis_final'][3]) <= 1e-10
    g = load_golden('evalpts_eul_shu_k2_nc100')
    e = Euler(make_parser().parse_args(['-ncell', '100', '-degree', '2', '-ic', 'shuosher', '-Tf', '0.2', '-char_lim', '1', '-tvbM', '100']))
    e.run()
    x, r, v, pre, lim = e.eval_points()
    assert np.array_equal(x, g['x'])
    for got, key in ((r, 'r_final'), (v, 'v_final'), (pre, 'p_final')):
        assert rel_linf(got, g[key]) <= 1e-10
    assert np.allclose(lim[:, 0] - 0.05, g['ylim_final'][:, 0], rtol=0, atol=1e-9) and np.allclose(lim[:, 1] + 0.05, g['ylim_final'][:, 1], rtol=0, atol=1e-9)
