"""class Euler + command line: the drop-in for the script dg1d/python/euler/euler.py.

The reference is a module-level script; its surface is the flag set of euler.py:16-27 and
the globals rho1, mom1, ene1, it, t, dx, cfl, Mdx2.  Euler(args) mirrors class DG of the scalar solver
and keeps those names as attributes.  One extension: -num_flux; the default 'rusanov' is the
reference's only flux (gas.py:79-89); 'roe' restates dg2d/models/euler/pde.h:209-285; 'lf', 'ecusp',
'hllc', 'ausmdv', 'lfc' are the flux family of the C sibling (dg1d/c/euler/src/flux.c, selected there by
the 'flux' line of inp.dat, dg.h:7-11), bit-identical to that file in the reference loop.
Copy ("neumann") boundary faces, limiter on cells 1..nc-2 whenever degree > 0 (euler.py:239-242).
"""
import argparse

import numpy as np

from ._lib import EULER_FLUXES
from . import ic as ic_mod
from .device import Dg1dDevice


class Euler:
    def __init__(self, args, ic_plugin=None):
        if ic_plugin is None:
            if args.ic not in ic_mod.EULER:
                print('Unknown initial condition')           # euler.py:37-39
                raise SystemExit
            ic_plugin = ic_mod.EULER[args.ic]
        self.k = args.degree
        self.cfl = args.cfl / (2 * self.k + 1)               # euler.py:42
        self.nc = args.ncell
        self.nd = self.k + 1
        self.xmin, self.xmax = ic_plugin.xmin, ic_plugin.xmax
        self.ic = ic_plugin.initial_condition
        self.dx = (self.xmax - self.xmin) / self.nc
        self.Mdx2 = args.tvbM * self.dx ** 2                 # euler.py:47
        self.char_lim = args.char_lim
        self.Tf = args.Tf if args.Tf > 0.0 else ic_plugin.Tf  # euler.py:177-178
        self.num_flux = getattr(args, 'num_flux', 'rusanov')
        if self.num_flux not in EULER_FLUXES:
            print('Unknown numerical flux')
            raise SystemExit
        self.dev = None
        self.it, self.t = 0, 0.0

    def setup(self):
        self.dev = Dg1dDevice('euler', self.num_flux, self.k, self.nc, limiter=(self.k > 0),
                              char_lim=self.char_lim, bc=('copy', 'copy'))
        o = self.dev.ops
        self.Vf, self.Vg, self.Vu, self.bm, self.bp = o.Vf, o.Vg, o.Vu, o.bm, o.bp
        self.xg, self.wg, self.xu = o.xg, o.wg, o.xu
        self._res = self.dev.new_state()

    def set_ic(self):
        """euler.py:162-169."""
        x = ic_mod.gauss_points(self.xmin, self.dx, self.nc, self.xg)
        rho, mom, ene = self.ic(x.reshape(-1))
        vals = np.stack([np.asarray(rho, dtype=np.float64), np.asarray(mom, dtype=np.float64),
                         np.asarray(ene, dtype=np.float64)]).reshape(3, 1, self.nc, self.nd)
        self.dev.project(np.ascontiguousarray(vals))

    @property
    def u1(self):
        """(3, nc, nd): rho, mom, ene."""
        return self.dev.unpack().cpu().numpy().reshape(3, self.nc, self.nd)

    @u1.setter
    def u1(self, a):
        self.dev.pack(np.asarray(a, dtype=np.float64).reshape(3, 1, self.nc, self.nd))

    @property
    def u0(self):
        """(3, nc, nd): the state at the start of the last time step (rho0, mom0, ene0 of euler.py:187-189); zeros
        before the first step (euler.py:151-159)."""
        if self.it == 0:
            return np.zeros((3, self.nc, self.nd))
        return self.dev.unpack(self.dev.bufs[1]).cpu().numpy().reshape(3, self.nc, self.nd)

    rho0 = property(lambda self: self.u0[0])
    mom0 = property(lambda self: self.u0[1])
    ene0 = property(lambda self: self.u0[2])
    rho1 = property(lambda self: self.u1[0])
    mom1 = property(lambda self: self.u1[1])
    ene1 = property(lambda self: self.u1[2])

    @property
    def res(self):
        return self.dev.unpack(self._res).cpu().numpy().reshape(3, self.nc, self.nd)

    def compute_residual(self):
        self.dev.residual(self.dev.bufs[0], self._res)

    def apply_limiter(self):
        if self.k > 0:
            d = self.dev
            d.limiter(d.bufs[0], d.bufs[1], self.Mdx2)
            d.bufs[0], d.bufs[1] = d.bufs[1], d.bufs[0]

    def max_speed(self):
        return float(self.dev.max_speed().cpu()[0])

    def eval_points(self):
        """What euler.py:77-148 plots, headless: returns (x, r, v, pre, limits) -- density, velocity m/r and pressure at
        the uniform points of every cell, each (nc, nu), and limits[q] = (min, max) of quantity q (the reference sets
        its y-axes to min - 0.05, max + 0.05).  Computed on the GPU."""
        out, mm = self.dev.eval_points()
        i = np.arange(self.nc)
        xc = self.xmin + i * self.dx + 0.5 * self.dx
        x = xc[:, None] + (0.5 * self.dx * self.xu)[None, :]
        o = out.cpu().numpy().reshape(3, self.nc, -1)
        return x, o[0], o[1], o[2], mm.cpu().numpy().reshape(3, 2)

    def solve(self):
        """euler.py:180-274 as a device-resident loop of fused stages."""
        d = self.dev
        d.reset_control()
        it, t = d.solve(self.cfl, self.dx, self.Tf, self.Mdx2)
        self.it, self.t = int(it[0]), float(t[0])

    def run(self):
        self.setup()
        self.set_ic()
        self.solve()
        return self.it, self.t


def make_parser():
    p = argparse.ArgumentParser()
    p.add_argument('-ncell', type=int, help='Number of cells', default=100)
    p.add_argument('-degree', type=int, help='Polynomial degree', default=1)
    p.add_argument('-cfl', type=float, help='CFL number', default=0.9)
    p.add_argument('-Tf', type=float, help='Final time', default=0.0)
    p.add_argument('-plot_freq', type=int, help='Accepted for compatibility; plotting is not built', default=5)
    p.add_argument('-ic', choices=('sod', 'lax', 'shuosher'), help='Initial condition', default='sod')
    p.add_argument('-char_lim', type=int, help='Characteristic limiter', default=0)
    p.add_argument('-tvbM', type=float, help='TVB M parameter', default=0.0)
    p.add_argument('-num_flux', choices=EULER_FLUXES, default='rusanov',
                   help='Numerical flux (extension; rusanov is the reference; lf..lfc: dg1d/c/euler/src/flux.c)')
    return p


def main(argv=None):
    args = make_parser().parse_args(argv)
    e = Euler(args)
    it, t = e.run()
    u = e.u1
    print('%d cells, Deg = %d, CFL = %s, it = %d, t = %.3e' % (e.nc, e.k, e.cfl, it, t))
    print('mass, momentum, energy = %.16e %.16e %.16e' % tuple(u[:, :, 0].sum(axis=1) * e.dx))
    return e


if __name__ == '__main__':
    main()
