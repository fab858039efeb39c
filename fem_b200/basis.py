"""Orthonormal-in-L2 Legendre modal basis phi_n = sqrt(2n+1) P_n on [-1,1] and the small operators
of the DG scheme.

Interface follows dg1d/python/scalar/basis.py:4-39 (Legendre, dLegendre, shape_value, shape_grad,
Vandermonde) so the class DG can expose the same helpers.  Internally every polynomial of degree
<= k is produced in one pass of the three-term recurrence; the arithmetic per step is the
reference's expression, so the resulting doubles are identical (tests compare bit for bit with
the operator arrays stored in the goldens).  SURVEY F6: these constants are uploaded to the GPU,
never re-derived there.
"""
import numpy as np


def _tables(k, x):
    """Values P_0..P_k and derivatives P'_0..P'_k at x (scalar or array)."""
    x = np.asarray(x, dtype=np.float64)
    val = [np.ones_like(x)]
    der = [np.zeros_like(x)]
    if k >= 1:
        val.append(x + 0.0)
        der.append(np.ones_like(x))
    for n in range(2, k + 1):
        val.append((2.0 * n - 1.0) / n * x * val[n - 1] - (n - 1.0) / n * val[n - 2])
        der.append(n * val[n - 1] + x * der[n - 1])
    return val, der


def Legendre(n, x):
    return _tables(n, x)[0][n]


def dLegendre(n, x):
    return _tables(n, x)[1][n]


def shape_value(n, x):
    return Legendre(n, x) * np.sqrt(2 * n + 1)


def shape_grad(n, x):
    return dLegendre(n, x) * np.sqrt(2 * n + 1)


def Vandermonde(k, x):
    """V[i, j] = phi_j(x_i)."""
    x = np.asarray(x, dtype=np.float64)
    val, _ = _tables(k, x)
    V = np.empty((len(x), k + 1))
    for j in range(k + 1):
        V[:, j] = val[j] * np.sqrt(2 * j + 1)
    return V


class Operators:
    """xg, wg, Vf, Vg, Vu, xu, bm, bp as set up in dgsolver.py:67-99 / euler.py:50-74."""

    def __init__(self, k):
        nd = k + 1
        self.k = k
        self.xg, self.wg = np.polynomial.legendre.leggauss(nd)
        self.Vf = Vandermonde(k, self.xg)
        _, der = _tables(k, self.xg)
        # Vg[i, j] = phi_i'(xg_j) * wg_j  (row = test function, column = quadrature point)
        self.Vg = np.empty((nd, nd))
        for i in range(nd):
            self.Vg[i, :] = der[i] * np.sqrt(2 * i + 1) * self.wg
        nu = max(2, k + 1)
        self.xu = np.linspace(-1.0, +1.0, nu)
        self.Vu = Vandermonde(k, self.xu)
        self.bm = Vandermonde(k, np.array([-1.0]))[0].copy()
        self.bp = Vandermonde(k, np.array([+1.0]))[0].copy()


# SSP-RK3 coefficients, dgsolver.py:7-8 / euler.py:8-9.  brk is 1.0 - ark, NOT the literal 2/3 (F6).
ark = np.array([0.0, 3.0 / 4.0, 1.0 / 3.0])
brk = 1.0 - ark
sqrt3 = np.sqrt(3.0)
isqrt3 = 1.0 / sqrt3
