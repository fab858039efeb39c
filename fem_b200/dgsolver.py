"""class DG: scalar conservation laws with periodic BC, the drop-in for dg1d/python/scalar/dgsolver.py.
Extension: args.bc = 'dirichlet' closes the mesh as the C++ twin does (dg1d/scalar_legendre/dg.h:367-413,
README.md:91-137): inflow on the left with boundary_condition(x, t), outflow on the right; default 'periodic'.

Same constructor namespace (degree, cfl, ncell, Tf, plot_freq, limit, compute_error, ic, tvbM, pde,
num_flux), same method names (setup, set_ic, compute_residual, apply_limiter, compute_error_norm,
solve, run) and same attributes (k, nd, nc, dx, cfl, Mdx2, xmin, xmax, Vf, Vg, Vu, bm, bp, xg, wg, xu,
u0, u1, res) as dgsolver.py:10-294.  The arrays u0/u1/res live on the GPU in SoA form; the
attributes return (nc, nd) numpy copies and assigning to `u1` uploads.  Plotting is left out
(no display on the target), so plot_freq is accepted and ignored.
PDE 'varadv' evaluates a(x)^5 by repeated squaring where numpy calls pow(): parity for it is 1e-15-level, not bitwise.
"""
import numpy as np

from . import ic as ic_mod
from .basis import Vandermonde
from .device import Dg1dDevice, resolve_scalar_flux


class DG:
    def __init__(self, args):
        self.k = args.degree
        self.cfl = args.cfl / (2 * args.degree + 1)          # dgsolver.py:13
        self.nc = args.ncell
        self.Tf = args.Tf
        self.plot_freq = args.plot_freq
        self.limit = args.limit
        self.compute_error = args.compute_error
        self.nd = args.degree + 1
        if args.ic not in ic_mod.SCALAR:
            print('Unknown initial condition')               # dgsolver.py:33-35
            raise SystemExit
        plug = ic_mod.SCALAR[args.ic]
        self.xmin, self.xmax = plug.xmin, plug.xmax
        self.ic = plug.initial_condition
        self.dx = (self.xmax - self.xmin) / self.nc
        self.Mdx2 = args.tvbM * self.dx ** 2                 # dgsolver.py:40
        if args.pde not in ('linear', 'varadv', 'burger'):
            print('PDE not implemented')                     # dgsolver.py:60-62
            raise SystemExit
        self.pde = args.pde
        self.num_flux = resolve_scalar_flux(args.pde, args.num_flux)
        self.bc = getattr(args, 'bc', 'periodic')
        if self.bc not in ('periodic', 'dirichlet'):
            print('Unknown boundary condition')
            raise SystemExit
        # the twin's exercise (README.md:91-137): u(xmin, t) = exact solution there; for unit-speed advection that is
        # ic(x - t).  For the other PDEs the default holds the initial boundary value; replace the attribute to change it.
        if self.pde == 'linear':
            self.boundary_condition = lambda x, t: float(self.ic(np.array([x - t]))[0])
        else:
            self.boundary_condition = lambda x, t: float(self.ic(np.array([x]))[0])
        self.dev = None
        self.it, self.t = 0, 0.0

    # ---- set-up -----------------------------------------------------------------------------------------
    def setup(self):
        self.dev = Dg1dDevice(self.pde, self.num_flux, self.k, self.nc, limiter=(self.limit == 'yes'),
                              bc=('periodic', 'periodic') if self.bc == 'periodic' else ('inflow', 'outflow'))
        self.dev.set_mesh(self.xmin, self.dx)
        o = self.dev.ops
        self.Vf, self.Vg, self.Vu, self.bm, self.bp = o.Vf, o.Vg, o.Vu, o.bm, o.bp
        self.xg, self.wg, self.xu = o.xg, o.wg, o.xu
        self._res = self.dev.new_state()

    def set_ic(self):
        """L2 projection of the initial condition, dgsolver.py:107-122: the IC is evaluated on the host
        in one vectorised call, the projection runs on the GPU."""
        x = ic_mod.gauss_points(self.xmin, self.dx, self.nc, self.xg)
        vals = np.ascontiguousarray(self.ic(x.reshape(-1)).reshape(1, 1, self.nc, self.nd))
        self.dev.project(vals)

    # ---- reference-layout views -----------------------------------------------------------------------------
    @property
    def u1(self):
        return self.dev.unpack().cpu().numpy().reshape(self.nc, self.nd)

    @u1.setter
    def u1(self, a):
        self.dev.pack(np.asarray(a, dtype=np.float64).reshape(1, 1, self.nc, self.nd))

    @property
    def u0(self):
        """The state at the start of the last time step (dgsolver.py:269 copies u1 into it at every step); zeros before
        the first step, as dgsolver.py:102 allocates it.  On the device it is the second of the three rotating buffers."""
        if self.it == 0:
            return np.zeros((self.nc, self.nd))
        return self.dev.unpack(self.dev.bufs[1]).cpu().numpy().reshape(self.nc, self.nd)

    @property
    def res(self):
        return self.dev.unpack(self._res).cpu().numpy().reshape(self.nc, self.nd)

    # ---- the unfused pieces (same names as the reference) ------------------------------------------------
    def compute_residual(self):
        self.dev.residual(self.dev.bufs[0], self._res)

    def apply_limiter(self):
        if self.limit == 'yes':
            d = self.dev
            d.limiter(d.bufs[0], d.bufs[1], self.Mdx2)
            d.bufs[0], d.bufs[1] = d.bufs[1], d.bufs[0]

    def max_speed(self):
        return float(self.dev.max_speed().cpu()[0])

    # ---- error norm, dgsolver.py:232-252: exact solution evaluated on the host, quadrature and sum on the device ----
    def compute_error_norm(self):
        k, nc, xmin, dx = self.k, self.nc, self.xmin, self.dx
        ng = k + 3
        xg, wg = np.polynomial.legendre.leggauss(ng)
        V = Vandermonde(k, xg)
        x = ic_mod.gauss_points(xmin, dx, nc, xg)
        if self.bc == 'dirichlet' and self.pde == 'linear':
            x = x - self.t          # exact at any time with the inflow data above (the periodic case relies on whole periods)
        ue = np.ascontiguousarray(self.ic(x.reshape(-1)).reshape(1, nc, ng))
        e2 = float(self.dev.error_norm(ue, V, wg, dx).cpu()[0])
        return nc, dx, np.sqrt(e2)

    # ---- what the reference plots (dgsolver.py:125-167), headless: values at the uniform points and the axis limits ----
    def eval_points(self):
        """Returns (x, f, (umin, umax)): x[i,:] = xc + 0.5*dx*xu and f[i,:] = Vu @ u1[i,:] for every cell, both (nc, nu),
        and the running extrema from which the reference sets its axis (umin - 0.1, umax + 0.1).  Computed on the GPU."""
        out, mm = self.dev.eval_points()
        i = np.arange(self.nc)
        xc = self.xmin + i * self.dx + 0.5 * self.dx
        x = xc[:, None] + (0.5 * self.dx * self.xu)[None, :]
        mm = mm.cpu().numpy()
        return x, out.cpu().numpy().reshape(self.nc, -1), (float(mm[0, 0, 0]), float(mm[0, 0, 1]))

    # ---- time loop ---------------------------------------------------------------------------------------------
    def solve(self):
        """dgsolver.py:255-278 as a device-resident loop of fused stages."""
        d = self.dev
        d.reset_control()
        if self.bc == 'dirichlet':
            side_x = {'left': self.xmin, 'right': self.xmax}
            speed = 1.0 if self.pde == 'linear' else d.speed_const if self.pde == 'varadv' else None
            self.it, self.t = d.solve_closed(self.cfl, self.dx, self.Tf, self.Mdx2,
                                             lambda side, t: self.boundary_condition(side_x[side], t), const_speed=speed)
            return
        it, t = d.solve(self.cfl, self.dx, self.Tf, self.Mdx2)
        self.it, self.t = int(it[0]), float(t[0])

    def run(self):
        self.setup()
        self.set_ic()
        self.solve()
        if self.compute_error == 'yes':
            return self.compute_error_norm()
