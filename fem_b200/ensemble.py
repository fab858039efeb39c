"""Batched ensemble of independent 1-D Euler Riemann problems (BASELINE config 5): one kernel launch per RK stage
advances every problem, each with its own dt (euler.py:182 evaluated per problem on the device); problems that
reach Tf are passed through.  Across GPUs the problems are sharded and nothing is communicated.

Each member is what `python euler.py -ic sod` computes with sod.py's states replaced through
riemann.riemann_data(vl, vr, xs, x) (riemann.py:4-13): vl, vr = (rho, vel, pre) left/right of xs on [xmin, xmax]."""
import numpy as np

from . import ic as ic_mod
from .device import Dg1dDevice
from .distributed import shard_problems


class EulerEnsemble:
    def __init__(self, vl, vr, xs=0.5, ncell=100, degree=1, cfl=0.9, Tf=0.2, char_lim=0, tvbM=0.0, num_flux='rusanov',
                 xmin=0.0, xmax=1.0, world=1, rank=0):
        vl, vr = np.atleast_2d(np.asarray(vl, dtype=np.float64)), np.atleast_2d(np.asarray(vr, dtype=np.float64))
        assert vl.shape == vr.shape and vl.shape[1] == 3
        self.first, self.last = shard_problems(vl.shape[0], world, rank)     # this rank's problems
        self.vl, self.vr = vl[self.first:self.last], vr[self.first:self.last]
        self.xs = np.broadcast_to(np.asarray(xs, dtype=np.float64), (vl.shape[0],))[self.first:self.last]
        self.nprob = self.last - self.first
        self.k, self.nd, self.nc = degree, degree + 1, ncell
        self.cfl = cfl / (2 * degree + 1)                    # euler.py:42
        self.xmin, self.xmax = xmin, xmax
        self.dx = (xmax - xmin) / ncell
        self.Mdx2 = tvbM * self.dx ** 2                      # euler.py:47
        self.Tf = Tf
        self.dev = Dg1dDevice('euler', num_flux, degree, ncell, limiter=(degree > 0), char_lim=char_lim, bc=('copy', 'copy'),
                              nprob=self.nprob)
        self.it = self.t = None

    def set_ic(self):
        x = ic_mod.gauss_points(self.xmin, self.dx, self.nc, self.dev.ops.xg)
        vals = np.empty((3, self.nprob, self.nc, self.nd))
        for p in range(self.nprob):
            vals[0, p], vals[1, p], vals[2, p] = ic_mod.riemann_data(self.vl[p], self.vr[p], self.xs[p], x)
        self.dev.project(vals)

    def solve(self):
        self.dev.reset_control()
        self.it, self.t = self.dev.solve(self.cfl, self.dx, self.Tf, self.Mdx2)
        return self.it, self.t

    def run(self):
        self.set_ic()
        return self.solve()

    @property
    def u1(self):
        """(nprob, 3, nc, nd): rho, mom, ene coefficients of every problem of this rank."""
        return self.dev.unpack().cpu().numpy().transpose(1, 0, 2, 3)
