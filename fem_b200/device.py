"""Dg1dDevice: device-resident state + the calls into libdg1d.so.

PyTorch supplies device memory, streams and (in distributed.py) the process group; every kernel is
in libdg1d.so and is reached through ctypes with raw pointers.  There is no CPU path: constructing a
Dg1dDevice without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from .basis import Operators, ark, brk, isqrt3, sqrt3


def resolve_scalar_flux(pde, num_flux):
    """Numerical-flux selection of dgsolver.py:43-59."""
    if pde == 'varadv':
        return 'upwind'          # dgsolver.py:49-51: varadv.py brings its own (upwind) numflux
    if pde == 'linear':
        return 'central' if num_flux == 'central' else 'upwind'
    if pde == 'burger':
        if num_flux == 'central':
            return 'central'
        if num_flux in ('roe', 'upwind'):
            return 'roe'
        return 'godunov'
    raise ValueError('PDE not implemented')


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Dg1dDevice:
    """One block of cells (x nprob independent problems) resident on one GPU.

    State lives in three rotating SoA buffers (u^n, and the two stage buffers); bufs[0] is always
    the current solution.  Per-problem control scalars (t, dt, lam, speed, it, done) live on the
    device too, so the time loop runs without host round trips (dg1d_run_steps).
    """

    def __init__(self, pde, flux, degree, ncell, limiter=False, char_lim=0, bc=('periodic', 'periodic'), nprob=1,
                 cell_offset=0, ncell_global=None, gamma=1.4, device=None):
        if not torch.cuda.is_available():
            raise _lib.Dg1dError('no CUDA device: fem_b200 has no CPU fallback')
        self.lib = _lib.load()
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.pde, self.flux, self.k, self.nd = pde, flux, int(degree), int(degree) + 1
        self.nc, self.nprob = int(ncell), int(nprob)
        self.nvar = 3 if pde == 'euler' else 1
        cfg = _lib.Config(pde=_lib.PDE[pde], flux=_lib.FLUX[flux], degree=self.k,
                          limiter=_lib.LIMITER['tvb' if limiter else 'none'], char_lim=int(char_lim),
                          bc_left=_lib.BC[bc[0]], bc_right=_lib.BC[bc[1]], nprob=self.nprob, ncell=self.nc,
                          cell_offset=int(cell_offset),
                          ncell_global=int(self.nc if ncell_global is None else ncell_global), gamma=float(gamma))
        self.bc = tuple(bc)
        h = C.c_void_p()
        _lib.check(self.lib.dg1d_create(C.byref(cfg), C.byref(h)))
        self.h = h
        self.ops = Operators(self.k)
        o = self.ops
        _lib.check(self.lib.dg1d_set_operators(
            self.h, _dptr(np.ascontiguousarray(o.Vf)), _dptr(np.ascontiguousarray(o.Vg)), _dptr(o.bm), _dptr(o.bp),
            _dptr(np.ascontiguousarray(o.wg)), _dptr(ark), _dptr(brk), float(sqrt3), float(isqrt3)))
        self.pitch = int(self.lib.dg1d_pitch(self.h))
        self.state_doubles = int(self.lib.dg1d_state_doubles(self.h))
        with torch.cuda.device(self.device):
            self.bufs = [torch.zeros(self.state_doubles, dtype=torch.float64, device=self.device) for _ in range(3)]
            # control block, one entry per problem
            self.t = torch.zeros(self.nprob, dtype=torch.float64, device=self.device)
            self.dt = torch.zeros_like(self.t)
            self.lam = torch.zeros_like(self.t)
            self.speed = torch.zeros_like(self.t)
            self.it = torch.zeros(self.nprob, dtype=torch.int64, device=self.device)
            self.done = torch.zeros(self.nprob, dtype=torch.int32, device=self.device)
            self.dt_log = None
        self.aos_shape = (self.nvar, self.nprob, self.nc, self.nd)

    def set_mesh(self, xmin, dx):
        """Mesh origin and cell size for x-dependent fluxes (pde 'varadv'); max_speed as varadv.py:15-17 evaluates it."""
        x = np.linspace(-1.0, 1.0, 100)
        speed = float(np.abs(1.0 + (1.0 - x ** 2) ** 5).max())
        self.speed_const = speed
        _lib.check(self.lib.dg1d_set_mesh(self.h, float(xmin), float(dx), _dptr(np.ascontiguousarray(self.ops.xg)), speed))

    def __del__(self):
        h = getattr(self, 'h', None)
        if h is not None and h.value:
            self.lib.dg1d_destroy(h)
            self.h = None

    # ---- plumbing -------------------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _to_dev(self, a):
        if isinstance(a, torch.Tensor):
            return a.to(device=self.device, dtype=torch.float64).contiguous()
        a = np.ascontiguousarray(a, dtype=np.float64)
        return torch.from_numpy(a).to(self.device, non_blocking=False)

    def new_state(self):
        return torch.zeros(self.state_doubles, dtype=torch.float64, device=self.device)

    def reset_control(self):
        for x in (self.t, self.dt, self.lam, self.speed, self.it, self.done):
            x.zero_()

    # ---- state in / out ------------------------------------------------------------------------------------
    def pack(self, aos, out=None):
        """aos (nvar, nprob, nc, nd), host or device -> SoA state buffer (default: the current one)."""
        aos = self._to_dev(aos)
        assert aos.numel() == int(np.prod(self.aos_shape)), (tuple(aos.shape), self.aos_shape)
        out = self.bufs[0] if out is None else out
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_pack_state(self.h, aos.data_ptr(), out.data_ptr(), self._stream()))
        return out

    def unpack(self, soa=None, out=None):
        """SoA state buffer (default: current) -> device tensor (nvar, nprob, nc, nd); `out`: write into this tensor
        (e.g. staging()) instead of allocating a new one."""
        soa = self.bufs[0] if soa is None else soa
        aos = torch.empty(self.aos_shape, dtype=torch.float64, device=self.device) if out is None else out
        assert aos.is_contiguous() and aos.numel() == int(np.prod(self.aos_shape)) and aos.dtype == torch.float64
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_unpack_state(self.h, soa.data_ptr(), aos.data_ptr(), self._stream()))
        return aos

    # ---- host state in and out, streamed ---------------------------------------------------------------------
    STAGE_BYTES = int(os.environ.get('DG1D_STAGE_MB', '32')) << 20      # per staging slot; two slots

    def _stream_plan(self):
        """Chunks of cells and the two staging slots (kept for the life of the object: moving a state allocates
        nothing after the first call).  One problem per handle: the host array is one contiguous (nc, nd) block per
        variable (the reference's u1 / rho1, mom1, ene1), so a chunk is nvar copies.  Ensembles go in one piece."""
        if getattr(self, '_plan', None) is None:
            per_cell = self.nvar * self.nprob * self.nd * 8
            n = self.nc if self.nprob > 1 else max(1, min(self.nc, self.STAGE_BYTES // per_cell))
            slots = [torch.empty((self.nvar, self.nprob, n, self.nd), dtype=torch.float64, device=self.device) for _ in range(2)]
            copy_stream = torch.cuda.Stream(self.device)
            free = [torch.cuda.Event() for _ in range(2)]       # slot may be overwritten
            full = [torch.cuda.Event() for _ in range(2)]       # slot holds the chunk
            self._plan = (n, slots, copy_stream, free, full)
        return self._plan

    def upload(self, host_aos):
        """Host state in the reference layout (nvar, nprob, nc, nd) -- numpy, or a torch tensor in pinned memory for
        asynchronous copies -- into the current SoA buffer.  Streamed: the host->device copy of one chunk (copy stream)
        overlaps the layout conversion of the one before (dg1d_pack_cells, current stream); two 32 MB staging slots
        instead of a second full-size array on the device."""
        t = host_aos if isinstance(host_aos, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(host_aos, dtype=np.float64))
        t = t.reshape(self.aos_shape)
        n, slots, cs, free, full = self._stream_plan()
        cur = torch.cuda.current_stream(self.device)
        out = self.bufs[0]
        with torch.cuda.device(self.device):
            for k, c0 in enumerate(range(0, self.nc, n)):
                m, sl = min(n, self.nc - c0), k % 2
                cs.wait_event(free[sl]) if k >= 2 else cs.wait_stream(cur)
                with torch.cuda.stream(cs):
                    dst = slots[sl] if m == n else slots[sl].view(-1)[: self.nvar * self.nprob * m * self.nd].view(self.nvar, self.nprob, m, self.nd)
                    if m == self.nc:
                        dst.copy_(t, non_blocking=True)
                    else:                 # one contiguous (m, nd) block per variable (chunking implies nprob == 1)
                        for v in range(self.nvar):
                            dst[v, 0].copy_(t[v, 0, c0:c0 + m], non_blocking=True)
                    full[sl].record(cs)
                cur.wait_event(full[sl])
                _lib.check(self.lib.dg1d_pack_cells(self.h, dst.data_ptr(), out.data_ptr(), c0, m, self._stream()))
                free[sl].record(cur)
        return out

    def download(self, host_out):
        """Current state -> host tensor in the reference layout (pinned memory makes the copies asynchronous; the call
        returns when they are queued: synchronise the device before reading).  Streamed like upload()."""
        t = host_out.reshape(self.aos_shape)
        n, slots, cs, free, full = self._stream_plan()
        cur = torch.cuda.current_stream(self.device)
        soa = self.bufs[0]
        with torch.cuda.device(self.device):
            for k, c0 in enumerate(range(0, self.nc, n)):
                m, sl = min(n, self.nc - c0), k % 2
                if k >= 2:
                    cur.wait_event(free[sl])
                src = slots[sl] if m == n else slots[sl].view(-1)[: self.nvar * self.nprob * m * self.nd].view(self.nvar, self.nprob, m, self.nd)
                _lib.check(self.lib.dg1d_unpack_cells(self.h, soa.data_ptr(), src.data_ptr(), c0, m, self._stream()))
                full[sl].record(cur)
                cs.wait_event(full[sl])
                with torch.cuda.stream(cs):
                    if m == self.nc:
                        t.copy_(src, non_blocking=True)
                    else:
                        for v in range(self.nvar):
                            t[v, 0, c0:c0 + m].copy_(src[v, 0], non_blocking=True)
                    free[sl].record(cs)
            cur.wait_stream(cs)       # work queued later on the current stream (and a synchronise on it) sees the copies done
        return host_out

    def project(self, vals, out=None):
        """IC point values at the Gauss points (nvar, nprob, nc, Nq) -> modal coefficients (a3)."""
        vals = self._to_dev(vals)
        assert vals.numel() == int(np.prod(self.aos_shape))
        out = self.bufs[0] if out is None else out
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_project(self.h, vals.data_ptr(), out.data_ptr(), self._stream()))
        return out

    # ---- kernels ----------------------------------------------------------------------------------------------
    def max_speed(self, soa=None):
        soa = self.bufs[0] if soa is None else soa
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_max_speed(self.h, soa.data_ptr(), self.speed.data_ptr(), self._stream()))
        return self.speed

    def step_control(self, cfl, dx, Tf, max_steps=-1):
        log_ptr, cap = (self.dt_log.data_ptr(), self.dt_log.shape[1]) if self.dt_log is not None else (None, 0)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_step_control(self.h, cfl, dx, Tf, max_steps, self.t.data_ptr(), self.dt.data_ptr(),
                                                  self.lam.data_ptr(), self.speed.data_ptr(), self.it.data_ptr(),
                                                  self.done.data_ptr(), log_ptr, cap, self._stream()))

    def stage(self, rk, u_in, u_n, u_out, Mdx2, with_speed=False, use_done=True, part=0):
        """part: 0 = every tile; 1 = the first and last tile of each problem; 2 = the others (dg1d_stage_part)."""
        if part:
            with torch.cuda.device(self.device):
                _lib.check(self.lib.dg1d_stage_part(self.h, part, rk, u_in.data_ptr(), None if u_n is None else u_n.data_ptr(),
                                                    u_out.data_ptr(), self.lam.data_ptr(),
                                                    self.done.data_ptr() if use_done else None, float(Mdx2),
                                                    self.speed.data_ptr() if with_speed else None, self._stream()))
            return
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_stage(self.h, rk, u_in.data_ptr(), None if u_n is None else u_n.data_ptr(),
                                           u_out.data_ptr(), self.lam.data_ptr(),
                                           self.done.data_ptr() if use_done else None, float(Mdx2),
                                           self.speed.data_ptr() if with_speed else None, self._stream()))

    def residual(self, u_in, res):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_residual(self.h, u_in.data_ptr(), res.data_ptr(), self._stream()))
        return res

    def rk_update(self, rk, u_in, u_n, res, u_out):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_rk_update(self.h, rk, u_in.data_ptr(), u_n.data_ptr(), res.data_ptr(),
                                               self.lam.data_ptr(), u_out.data_ptr(), self._stream()))
        return u_out

    def limiter(self, u_in, u_out, Mdx2):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_limiter(self.h, u_in.data_ptr(), u_out.data_ptr(), float(Mdx2), self._stream()))
        return u_out

    def enable_dt_log(self, cap):
        self.dt_log = torch.zeros((self.nprob, int(cap)), dtype=torch.float64, device=self.device)

    def run_steps(self, nsteps, cfl, dx, Tf, Mdx2, max_steps=-1):
        """nsteps x (step control + 3 fused stages), no host synchronisation.  Single-rank only."""
        ptrs = (C.c_void_p * 3)(*[b.data_ptr() for b in self.bufs])
        log_ptr, cap = (self.dt_log.data_ptr(), self.dt_log.shape[1]) if self.dt_log is not None else (None, 0)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_run_steps(self.h, int(nsteps), ptrs, cfl, dx, Tf, int(max_steps), float(Mdx2),
                                               self.t.data_ptr(), self.dt.data_ptr(), self.lam.data_ptr(),
                                               self.speed.data_ptr(), self.it.data_ptr(), self.done.data_ptr(),
                                               log_ptr, cap, self._stream()))
        by_ptr = {b.data_ptr(): b for b in self.bufs}
        self.bufs = [by_ptr[ptrs[i]] for i in range(3)]

    # ---- closed ends of the scalar PDEs: inflow / outflow (dg1d/scalar_legendre/dg.h:367-413) ----------
    def set_boundary(self, values, stride=0):
        """Prescribed {left, right} states of the inflow faces: one pair (stride 0) or a table with one pair per stage
        (`stride` doubles apart; dg1d_run_steps reads pair 3*s + rk).  The tensor is kept alive by this object."""
        self._bdry = None if values is None else self._to_dev(values).reshape(-1)
        _lib.check(self.lib.dg1d_set_boundary(self.h, None if self._bdry is None else self._bdry.data_ptr(), int(stride)))

    @staticmethod
    def stage_times(t, dt, ark, brk):
        """SSPRK3 stage times as the twin advances them (dg.h:539, :602): t, then ark*t + brk*(stage_time + dt)."""
        s0 = t
        s1 = ark[0] * t + brk[0] * (s0 + dt)
        s2 = ark[1] * t + brk[1] * (s1 + dt)
        return s0, s1, s2

    def solve_closed(self, cfl, dx, Tf, Mdx2, boundary, const_speed=None, max_steps=-1):
        """The time loop with inflow faces fed by boundary(side, t) -> state (side 'left' / 'right'), evaluated on the
        host at the stage times as the twin evaluates its Function.  const_speed (linear: 1.0; varadv: its constant):
        dt is known in advance, so the whole table is uploaded once and the loop runs without host round trips;
        otherwise (Burgers) dt depends on the solution and every step reads the wave speed back.  Returns (it, t)."""
        self.max_speed()
        t, it = float(self.t[0].item()), 0
        if const_speed is not None:
            rows, tt = [], t
            while tt < Tf and (max_steps < 0 or len(rows) < max_steps):      # the control kernel's arithmetic, dg1d_api.cu
                d = cfl * dx / const_speed
                if tt + d > Tf:
                    d = Tf - tt
                rows.append([[boundary('left', s), boundary('right', s)] for s in self.stage_times(tt, d, ark, brk)])
                tt = tt + d
            if rows:
                self.set_boundary(np.asarray(rows, dtype=np.float64), stride=2)
                self.run_steps(len(rows), cfl, dx, Tf, Mdx2, max_steps)
            return len(rows), float(self.t[0].item())
        while t < Tf and (max_steps < 0 or it < max_steps):
            d = cfl * dx / float(self.speed[0].item())
            if t + d > Tf:
                d = Tf - t
            self.set_boundary(np.asarray([[boundary('left', s), boundary('right', s)] for s in self.stage_times(t, d, ark, brk)]),
                              stride=2)
            self.run_steps(1, cfl, dx, Tf, Mdx2, -1)
            t, it = float(self.t[0].item()), it + 1
        return it, t

    def run_steps_graphed(self, nsteps, cfl, dx, Tf, Mdx2, max_steps=-1, unit=8):
        """run_steps through a CUDA graph of `unit` steps (even, so that the three rotating buffers are back in
        place after a replay), captured on first use and replayed afterwards.  Small meshes are launch-bound:
        a step is four kernels of a few microseconds each (config 1: 200 cells), and a graph replay costs one
        submission for 4*unit of them.  The first `unit` steps run eagerly, so that every kernel has been loaded
        before the capture begins."""
        assert unit % 2 == 0
        if not hasattr(self, '_graphs'):
            self._graphs, self._graph_warm = {}, False
        if nsteps < 2 * unit:
            return self.run_steps(nsteps, cfl, dx, Tf, Mdx2, max_steps)
        if not self._graph_warm:
            self.run_steps(unit, cfl, dx, Tf, Mdx2, max_steps)
            nsteps -= unit
            self._graph_warm = True
        key = (cfl, dx, Tf, float(Mdx2), int(max_steps), unit, tuple(b.data_ptr() for b in self.bufs),
               None if self.dt_log is None else self.dt_log.data_ptr())
        g = self._graphs.get(key)
        if g is None:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.graph(g, stream=side):
                self.run_steps(unit, cfl, dx, Tf, Mdx2, max_steps)     # recorded, not executed
            torch.cuda.current_stream(self.device).wait_stream(side)
            self._graphs[key] = g
        reps, rem = divmod(nsteps, unit)
        for _ in range(reps):
            g.replay()
        if rem:
            self.run_steps(rem, cfl, dx, Tf, Mdx2, max_steps)

    def solve(self, cfl, dx, Tf, Mdx2, max_steps=-1, chunk=16, max_chunk=512, graph=False):
        """The reference's `while t < Tf` loop (dgsolver.py:263-275 / euler.py:181-274), device-resident.
        The host only polls the done flags between chunks of steps.  Returns (it, t) as numpy arrays."""
        self.max_speed()
        if self.nprob == 1 and not graph:
            return self._solve_exact(cfl, dx, Tf, Mdx2, max_steps, max_chunk)
        while True:
            (self.run_steps_graphed if graph else self.run_steps)(chunk, cfl, dx, Tf, Mdx2, max_steps)
            if bool(self.done.all().item()):
                break
            # size the next chunk from the remaining time of the slowest problem (never relied on
            # for correctness: problems that reach Tf early are passed through by the kernels)
            t = self.t.cpu().numpy(); dt = self.dt.cpu().numpy()
            with np.errstate(all='ignore'):
                rem = np.nanmax(np.where(dt > 0, (Tf - t) / dt, 0.0))
            if max_steps >= 0:
                rem = min(rem, max_steps - int(self.it.min().item()))
            chunk = int(min(max(rem * 0.9, 1), max_chunk)) if np.isfinite(rem) else 1
        return self.it.cpu().numpy(), self.t.cpu().numpy()

    def _solve_exact(self, cfl, dx, Tf, Mdx2, max_steps, max_chunk):
        """One problem: the host tests `t < Tf` on the device's own t (dgsolver.py:263) and never launches a step the
        reference would not take, so that afterwards bufs[1] still holds the state at the start of the last step (the
        reference's u0, dgsolver.py:269; a passed-through step would overwrite it).  Chunks are at most half of the
        steps the current dt predicts, single steps over the last eight: dt would have to double inside a chunk to
        overshoot, and then only u0 is affected (finished problems are passed through by the kernels)."""
        launched = 0
        while True:
            t = float(self.t[0].item())
            if not (t < Tf) or (max_steps >= 0 and launched >= max_steps):
                break
            dt = float(self.dt[0].item())
            n = 1
            if launched > 0 and dt > 0.0 and np.isfinite(dt):
                rem = (Tf - t) / dt
                n = int(rem // 2) if rem >= 8.0 else 1
            n = max(1, min(n, max_chunk))
            if max_steps >= 0:
                n = min(n, max_steps - launched)
            self.run_steps(n, cfl, dx, Tf, Mdx2, max_steps)
            launched += n
        self.steps_launched = launched
        return self.it.cpu().numpy(), self.t.cpu().numpy()

    def error_norm(self, exact, V, w, dx, soa=None):
        """Sum over cells of 0.5*dx*sum_q (V[q,:].u - exact[:, i, q])^2 w[q] on the device (dgsolver.py:232-251 before
        the square root).  exact: (nprob, nc, ng) host or device.  Returns a device tensor [nprob]."""
        soa = self.bufs[0] if soa is None else soa
        V = np.ascontiguousarray(V, dtype=np.float64); w = np.ascontiguousarray(w, dtype=np.float64)
        ng = int(w.shape[0])
        exact = self._to_dev(exact)
        assert exact.numel() == self.nprob * self.nc * ng
        n = int(self.lib.dg1d_error_norm_scratch_doubles(self.h))
        scratch = torch.empty(n, dtype=torch.float64, device=self.device)
        out = torch.empty(self.nprob, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_error_norm(self.h, soa.data_ptr(), exact.data_ptr(), _dptr(V), _dptr(w), ng, float(dx),
                                                scratch.data_ptr(), out.data_ptr(), self._stream()))
        return out

    def eval_points(self, Vu=None, soa=None, with_minmax=True):
        """The solution at the uniform points of every cell, as the reference evaluates it for its plots
        (dgsolver.py:131-139; euler.py:83-100: density, velocity m/r, pressure), and the running (min, max) of each
        quantity.  Returns device tensors: values (nvar, nprob, nc, nu) and minmax (nvar, nprob, 2) or None."""
        soa = self.bufs[0] if soa is None else soa
        Vu = np.ascontiguousarray(self.ops.Vu if Vu is None else Vu, dtype=np.float64)
        nu = int(Vu.shape[0])
        out = torch.empty((self.nvar, self.nprob, self.nc, nu), dtype=torch.float64, device=self.device)
        mm = torch.empty((self.nvar, self.nprob, 2), dtype=torch.float64, device=self.device) if with_minmax else None
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_eval_points(self.h, soa.data_ptr(), _dptr(Vu), nu, out.data_ptr(),
                                                 None if mm is None else mm.data_ptr(), self._stream()))
        return out, mm

    # ---- halo (domain decomposition) ---------------------------------------------------------------------
    def halo_doubles(self):
        return int(self.lib.dg1d_halo_doubles(self.h))

    def halo_pack(self, soa, send_left, send_right):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_halo_pack(self.h, soa.data_ptr(),
                                               None if send_left is None else send_left.data_ptr(),
                                               None if send_right is None else send_right.data_ptr(), self._stream()))

    def halo_unpack(self, soa, recv_left, recv_right):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_halo_unpack(self.h, soa.data_ptr(),
                                                 None if recv_left is None else recv_left.data_ptr(),
                                                 None if recv_right is None else recv_right.data_ptr(), self._stream()))

    # ---- host-buffer entry point -------------------------------------------------------------------------
    def solve_host(self, u_aos, cfl, dx, Tf, Mdx2, max_steps=-1):
        """numpy in, numpy out through dg1d_solve_host (allocates its own device memory)."""
        u = np.ascontiguousarray(u_aos, dtype=np.float64).reshape(self.aos_shape).copy()
        t = np.zeros(self.nprob); it = np.zeros(self.nprob, dtype=np.int64)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dg1d_solve_host(self.h, _dptr(u), cfl, dx, Tf, int(max_steps), float(Mdx2), _dptr(t),
                                                it.ctypes.data_as(C.POINTER(C.c_int64))))
        return u, it, t
