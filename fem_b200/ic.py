"""Initial-condition plugins.  Same protocol as the reference's IC modules: an object exporting
xmin, xmax, initial_condition(x) (+ Tf for Euler) -- scalar: sin2pi.py, sin4pi.py, gauss.py,
hat.py, mult.py; Euler: sod.py, lax.py, shuosher.py, riemann.py.  All functions accept a 1-D (or
any-shape) array of points, so one vectorised call covers the whole mesh.
"""
import types

import numpy as np

gasGam = 1.4  # euler/gas.py:3


def _plug(name, xmin, xmax, fn, Tf=None):
    m = types.SimpleNamespace(name=name, xmin=xmin, xmax=xmax, initial_condition=fn)
    if Tf is not None:
        m.Tf = Tf
    return m


# ---- scalar -------------------------------------------------------------------------------------
def _hat(x):
    x = np.asarray(x)
    return np.where((x < 0.25) | (x > 0.75), 0.0, 1.0)


def _mult(x):  # Jiang-Shu composite wave, mult.py:2-16
    x = np.asarray(x, dtype=np.float64)
    f = np.zeros_like(x)
    m = (x > -0.8) & (x < -0.6)
    f[m] = np.exp(-np.log(2.0) * (x[m] + 0.7) ** 2 / 0.0009)
    m = (x > -0.4) & (x < -0.2)
    f[m] = 1.0
    m = (x > 0.0) & (x < 0.2)
    f[m] = 1.0 - np.abs(10.0 * (x[m] - 0.1))
    m = (x > 0.4) & (x < 0.6)
    f[m] = np.sqrt(1.0 - 100.0 * (x[m] - 0.5) ** 2)
    return f


SCALAR = {
    'sin2pi': _plug('sin2pi', 0.0, 1.0, lambda x: 1 + np.sin(2 * np.pi * x)),
    'sin4pi': _plug('sin4pi', -1.0, 1.0, lambda x: np.sin(4 * np.pi * x)),
    'gauss': _plug('gauss', 0.0, 1.0, lambda x: np.exp(-100 * (x - 0.5) ** 2)),
    'hat': _plug('hat', 0.0, 1.0, _hat),
    'mult': _plug('mult', -1.0, 1.0, _mult),
}


# ---- euler ------------------------------------------------------------------------------------------
def riemann_data(vl, vr, xs, x):
    """Two constant primitive states (rho, vel, pre) separated at xs; riemann.py:4-13."""
    rhol, vell, prel = vl
    rhor, velr, prer = vr
    enel = prel / (gasGam - 1.0) + 0.5 * rhol * vell ** 2
    ener = prer / (gasGam - 1.0) + 0.5 * rhor * velr ** 2
    left, right = (x <= xs), (x > xs)
    rho = left * rhol + right * rhor
    mom = left * rhol * vell + right * rhor * velr
    ene = left * enel + right * ener
    return rho, mom, ene


def _shuosher(x):  # shuosher.py:7-13
    eps = 0.2
    pre_shock = (x < -4.0)
    post = (x >= -4.0)
    vel = pre_shock * 2.629369
    pre = pre_shock * 10.33333 + post * 1.0
    rho = pre_shock * 3.857143 + post * (1 + eps * np.sin(5 * x))
    ene = pre / (gasGam - 1.0) + 0.5 * rho * vel ** 2
    return rho, rho * vel, ene


EULER = {
    'sod': _plug('sod', 0.0, 1.0, lambda x: riemann_data([1.0, 0.0, 1.0], [0.125, 0.0, 0.1], 0.5, x), Tf=0.2),
    'lax': _plug('lax', -5.0, 5.0, lambda x: riemann_data([0.445, 0.698, 3.528], [0.5, 0.0, 0.571], 0.0, x), Tf=1.3),
    'shuosher': _plug('shuosher', -5.0, 5.0, _shuosher, Tf=1.8),
}


def riemann_plugin(vl, vr, xs=0.5, xmin=0.0, xmax=1.0, Tf=0.2):
    """IC plug-in for an arbitrary Riemann problem (ensemble members)."""
    vl, vr = list(vl), list(vr)
    return _plug('riemann', xmin, xmax, lambda x: riemann_data(vl, vr, xs, x), Tf=Tf)


def gauss_points(xmin, dx, nc, xg):
    """x[i, q] = xc_i + 0.5 dx xg_q with xc_i = xmin + i dx + 0.5 dx (dgsolver.py:118-119)."""
    i = np.arange(nc)
    xc = xmin + i * dx + 0.5 * dx
    return xc[:, None] + (0.5 * dx * xg)[None, :]
