#!/usr/bin/env python3
"""bench.py -- DG DOF-updates/s per RK stage (BASELINE.json's metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c3|c2|c5] [--impl ours|reference]

A "step" is one SSP-RK3 time step (three fused stage kernels + the device-side step control) over the whole
mesh.  Default workload at every N is BASELINE config 4, weak-scaled: Euler Shu-Osher, degree 2, Rusanov flux,
characteristic TVB limiter (tvbM=100), 1.25e7 cells per GPU (1e8 cells at N=8), 1-D domain decomposition
with a two-cell halo per stage and one all_reduce(MAX) per step.  Each state buffer is 0.9 GB per GPU, far
larger than the 126 MB L2, so no L2 flush is needed between timed iterations.

value  : whole-job DOF-updates/s per RK stage = cells * (k+1) * nvar * 3K / T, state resident in HBM,
         T = CUDA-event time of the K steps, max over ranks.
e2e    : same metric through the public host-buffer API: pinned host state -> H2D -> pack -> K steps (the
         step's (t, dt) read back to the host after every step) -> unpack -> D2H, all inside the timed region.
roofline: HBM; algorithmic bytes = 64 B per DOF per step (stage 0: read u, write; stages 1,2: read u, read
         u^n, write) = 21.33 B per DOF-update, over the measured copy bandwidth.
cpu_baseline: the CPU oracle (a C port of the reference algorithm, OpenMP over all host cores) on a bounded
         sample of the same workload; the reference itself is single-threaded Python and does not travel.
--impl reference: only the CPU port is timed (rank 0); other ranks exit 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: pde, flux, degree, limiter, char_lim, tvbM, ic, periodic, cells per GPU, nprob per GPU, cfl
    'c4': dict(pde='euler', flux='rusanov', degree=2, limiter=True, char_lim=1, tvbM=100.0, ic='shuosher', periodic=False,
               cells=12_500_000, nprob=1, cfl=0.9,
               desc='config 4: Euler Shu-Osher, degree 2, Rusanov, characteristic TVB limiter (M=100), 1.25e7 cells/GPU, '
                    'domain-decomposed, 2-cell halo per stage'),
    'c4roe': dict(pde='euler', flux='roe', degree=2, limiter=True, char_lim=1, tvbM=100.0, ic='shuosher', periodic=False,
                  cells=12_500_000, nprob=1, cfl=0.9,
                  desc='config 4 as BASELINE.json names it: Euler Shu-Osher, degree 2, Roe flux (dg2d/models/euler/pde.h), characteristic '
                       'TVB limiter (M=100), 1.25e7 cells/GPU, domain-decomposed, 2-cell halo per stage'),
    'c3': dict(pde='burger', flux='roe', degree=2, limiter=True, char_lim=0, tvbM=10.0, ic='sin2pi', periodic=True,
               cells=10_000_000, nprob=1, cfl=0.9,
               desc='config 3: inviscid Burgers, degree 2, Roe flux, TVB limiter (M=10), 1e7 cells/GPU, periodic'),
    'c2': dict(pde='linear', flux='upwind', degree=1, limiter=False, char_lim=0, tvbM=0.0, ic='sin2pi', periodic=True,
               cells=1_000_000, nprob=1, cfl=0.9,
               desc='config 2: linear advection of a sine, upwind flux, no limiter, 1e6 cells/GPU (L2-resident)'),
    'c2big': dict(pde='linear', flux='upwind', degree=1, limiter=False, char_lim=0, tvbM=0.0, ic='sin2pi', periodic=True,
                  cells=20_000_000, nprob=1, cfl=0.9,
                  desc='config 2 beyond the L2: linear advection of a sine, upwind flux, no limiter, 2e7 cells/GPU'),
    'c5': dict(pde='euler', flux='rusanov', degree=1, limiter=True, char_lim=0, tvbM=0.0, ic='ensemble', periodic=False,
               cells=4096, nprob=512, cfl=0.9,
               desc='config 5: ensemble of independent Sod-type problems, degree 1, 4096 cells each, 512 problems/GPU, '
                    'per-problem dt, no communication'),
}


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        return float(json.load(open(p))['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


def gauss_points_range(xmin, dx, start, stop, xg):
    i = np.arange(start, stop)
    xc = xmin + i * dx + 0.5 * dx
    return xc[:, None] + (0.5 * dx * xg)[None, :]


def initial_values(w, k, xg, start, stop, ncg, prob0=0):
    """Point values of the IC at the Gauss points of local cells: (nvar, nprob, nc, Nq)."""
    from fem_b200 import ic as ic_mod
    nd = k + 1
    nc = stop - start
    if w['ic'] == 'ensemble':
        n = w['nprob']
        rng = np.random.default_rng(20260101)
        tot = 4096
        rl = rng.uniform(0.5, 2.0, tot); pl = rng.uniform(0.5, 2.0, tot)
        rr = rng.uniform(0.1, 0.5, tot); pr = rng.uniform(0.05, 0.5, tot)
        ul = rng.uniform(-0.2, 0.2, tot); ur = rng.uniform(-0.2, 0.2, tot)
        dx = 1.0 / ncg
        x = gauss_points_range(0.0, dx, 0, nc, xg)
        vals = np.empty((3, n, nc, nd))
        for p in range(n):
            q = (prob0 + p) % tot
            r, m, e = ic_mod.riemann_data([rl[q], ul[q], pl[q]], [rr[q], ur[q], pr[q]], 0.5, x)
            vals[0, p], vals[1, p], vals[2, p] = r, m, e
        return vals, dx, 0.2
    if w['pde'] == 'euler':
        plug = ic_mod.EULER[w['ic']]
        dx = (plug.xmax - plug.xmin) / ncg
        x = gauss_points_range(plug.xmin, dx, start, stop, xg)
        r, m, e = plug.initial_condition(x.reshape(-1))
        vals = np.stack([np.asarray(r, dtype=np.float64), np.asarray(m, dtype=np.float64), np.asarray(e, dtype=np.float64)])
        return vals.reshape(3, 1, nc, nd), dx, plug.Tf
    plug = ic_mod.SCALAR[w['ic']]
    dx = (plug.xmax - plug.xmin) / ncg
    x = gauss_points_range(plug.xmin, dx, start, stop, xg)
    return plug.initial_condition(x.reshape(-1)).reshape(1, 1, nc, nd), dx, 1.0e9


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
                                          '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons, pw = [], [], set(), []
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ts, ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 7:
                continue
            inside = (t0 - 0.05) <= ts <= (t1 + 0.15)
            try:
                if inside:
                    sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            if inside:
                for nme, val in zip(names, f[3:7]):
                    if val.lower().startswith('active'):
                        reasons.add(nme)
        if not sm:   # region shorter than the sampling period: use everything we saw
            for ts, ln in self.lines:
                f = [x.strip() for x in ln.split(',')]
                try:
                    sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
                except (ValueError, IndexError):
                    pass
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': float(np.max(mx)) if mx else None,
                'power_w_max': float(np.max(pw)) if pw else None, 'samples': len(sm), 'reasons': sorted(reasons)}


def post_shock_values(w, k, xg, start, stop, ncg):
    """Burgers sin2pi advanced to t = 0.25 (the shock forms at 1/(2 pi) = 0.159) on 2e4 cells with this solver, then
    evaluated at the Gauss points of fine cells [start, stop): the limiter is active around the shock."""
    import types
    from fem_b200.basis import Vandermonde
    from fem_b200.dgsolver import DG
    ncc = 20_000
    args = types.SimpleNamespace(pde='burger', ncell=ncc, nrefine=0, degree=k, cfl=w['cfl'], Tf=0.25, plot_freq=0, ic='sin2pi',
                                 limit='yes', tvbM=w['tvbM'], compute_error='no', num_flux='roe')
    dg = DG(args); dg.run()
    uc = dg.u1                                            # (ncc, nd)
    dx, dxc = 1.0 / ncg, 1.0 / ncc
    out = np.empty((1, 1, stop - start, k + 1))
    chunk = 2_000_000
    for a in range(start, stop, chunk):
        bnd = min(stop, a + chunk)
        x = gauss_points_range(0.0, dx, a, bnd, xg)       # (n, Nq)
        c = np.minimum((x / dxc).astype(np.int64), ncc - 1)
        xi = 2.0 * (x - (c + 0.5) * dxc) / dxc
        V = Vandermonde(k, xi.reshape(-1)).reshape(x.shape + (k + 1,))
        out[0, 0, a - start:bnd - start] = np.einsum('nqm,nqm->nq', V, uc[c])
    return out, dx, 1.0e9


def cpu_port(w, target_s=12.0, steps=None, warmup=0):
    """Times the CPU oracle (C port, OpenMP) on a bounded sample of workload `w`.  Returns (rate, info)."""
    from oracle import dg1d_oracle as orc
    from fem_b200.basis import Operators
    k = w['degree']
    nd = k + 1
    ncs = 200_000 if w['nprob'] == 1 else w['cells'] * 16
    cores = os.cpu_count() or 1
    ops = orc.make_ops(k)
    xg = Operators(k).xg
    if w['ic'] == 'ensemble':
        # 16 ensemble members back to back (the port has no batch dimension): time them one after another
        vals, dx, Tf = initial_values(dict(w, nprob=16), k, xg, 0, w['cells'], w['cells'])
        states = []
        for p in range(16):
            u = np.stack([orc.project(ops, Operators(k).wg, np.ascontiguousarray(vals[v, p])) for v in range(3)])
            states.append(np.ascontiguousarray(u))
        ndof = 16 * w['cells'] * nd * 3

        def run(n):
            for u in states:
                orc.euler_solve(ops, w['flux'], w['char_lim'], u, w['cfl'] / (2 * k + 1), dx, 1e9, w['tvbM'] * dx * dx, max_steps=n)
    else:
        vals, dx, Tf = initial_values(w, k, xg, 0, ncs, ncs)
        wg = Operators(k).wg
        nvar = vals.shape[0]
        if nvar == 3:
            u = np.ascontiguousarray(np.stack([orc.project(ops, wg, np.ascontiguousarray(vals[v, 0])) for v in range(3)]))
            run = lambda n: orc.euler_solve(ops, w['flux'], w['char_lim'], u, w['cfl'] / (2 * k + 1), dx, 1e9, w['tvbM'] * dx * dx,
                                            max_steps=n)
        else:
            u = orc.project(ops, wg, np.ascontiguousarray(vals[0, 0]))
            run = lambda n: orc.scalar_solve(ops, w['pde'], w['flux'], w['limiter'], u, w['cfl'] / (2 * k + 1), dx, 1e9,
                                             w['tvbM'] * dx * dx, max_steps=n)
        ndof = ncs * nd * nvar
    t0 = time.perf_counter(); run(1); t1 = time.perf_counter() - t0      # also warms the OpenMP team
    for _ in range(warmup):
        run(1)
    if steps is None:
        steps = int(max(2, min(60, target_s / max(t1, 1e-4))))
    t0 = time.perf_counter(); run(steps); dt = time.perf_counter() - t0
    rate = ndof * 3 * steps / dt
    return rate, dict(cores=cores, steps=steps, seconds=dt, ndof=ndof,
                      sample='%d cells x %d steps of the same workload (%s)' % (ndof // (nd * (3 if w['pde'] == 'euler' else 1)), steps,
                                                                                'C port of the reference algorithm, OpenMP'))


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on the first
    communicator): point file descriptor 1 at stderr for the duration of the run and keep the real stdout for the
    result line."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    return real


def _emit(real_stdout, obj):
    sys.stdout.flush()
    os.write(real_stdout, (json.dumps(obj) + '\n').encode())


SECONDARY = ('c4roe', 'c3', 'c5', 'c2', 'c2big')     # reported under "workloads" on the same JSON line
_VALS_CACHE = {}


def cpu_sample_desc(w):
    """The bounded sample of workload `w` that the CPU arms run (cpu_port below); stated in both arms' config."""
    if w['nprob'] == 1:
        return '200000 cells of the same mesh family, same scheme (the GPU arm runs %d cells per GPU)' % w['cells']
    return '16 of the ensemble members, %d cells each, one after another' % w['cells']


def make_config(name, w, world):
    """The config object of the JSON line: built from the workload alone, so both arms print the same keys and values."""
    k, nd = w['degree'], w['degree'] + 1
    nvar = 3 if w['pde'] == 'euler' else 1
    mb = w['cells'] * w['nprob'] * nd * nvar * 8 / 1e6
    return {'workload': w['desc'], 'name': name, 'degree': k, 'cells_per_gpu': w['cells'], 'problems_per_gpu': w['nprob'],
            'global_cells': w['cells'] * (world if w['nprob'] == 1 else 1), 'pde': w['pde'], 'flux': w['flux'],
            'limiter': 'tvb-char' if w['char_lim'] else ('tvb' if w['limiter'] else 'none'), 'tvbM': w['tvbM'],
            'l2_policy': ('state buffers (%.0f MB each, 3 rotating) exceed the 126 MB L2; no flush' % mb) if mb > 126 else
                         ('state buffers are %.0f MB each: L2-resident, a launch-latency figure rather than an HBM one' % mb),
            'cpu_reference_sample': cpu_sample_desc(w)}


class Case:
    """One workload resident on this rank's GPU: the rank's block (or its share of the ensemble), the IC projected."""

    def __init__(self, name, w, world, rank, transport='auto', native=True, post_shock=False, graph=False):
        import torch
        from fem_b200.distributed import DistributedDg1d
        from fem_b200.device import Dg1dDevice
        self.torch, self.name, self.w, self.world, self.rank, self.graph = torch, name, w, world, rank, graph
        k = w['degree']
        self.k, self.nd, self.nvar = k, k + 1, 3 if w['pde'] == 'euler' else 1
        self.state_note = None
        if w['nprob'] == 1:
            ncg = w['cells'] * world
            self.dd = DistributedDg1d(w['pde'], w['flux'], k, ncg, limiter=w['limiter'], char_lim=w['char_lim'], periodic=w['periodic'],
                                      native=native, transport=transport)
            self.dev = self.dd.dev
            key = (w['ic'], k, self.dd.start, self.dd.stop, ncg, post_shock)
            if key not in _VALS_CACHE:
                _VALS_CACHE.clear()          # one big array at a time
                if post_shock and w['pde'] == 'burger':
                    _VALS_CACHE[key] = post_shock_values(w, k, self.dev.ops.xg, self.dd.start, self.dd.stop, ncg)
                else:
                    _VALS_CACHE[key] = initial_values(w, k, self.dev.ops.xg, self.dd.start, self.dd.stop, ncg)
            if post_shock and w['pde'] == 'burger':
                self.state_note = 'Burgers advanced to t=0.25 (past shock formation) on 2e4 cells, re-projected'
            vals, dx, _ = _VALS_CACHE[key]
        else:   # ensemble: shard problems, no communication
            self.dd = None
            self.dev = Dg1dDevice(w['pde'], w['flux'], k, w['cells'], limiter=w['limiter'], char_lim=w['char_lim'], bc=('copy', 'copy'),
                                  nprob=w['nprob'])
            vals, dx, _ = initial_values(w, k, self.dev.ops.xg, 0, w['cells'], w['cells'], prob0=rank * w['nprob'])
        self.transport = self.dd.transport if (self.dd is not None and world > 1) else None
        self.dx, self.cfl, self.Mdx2 = dx, w['cfl'] / (2 * k + 1), w['tvbM'] * dx * dx
        self.Tf = 1.0e9       # throughput run: never reach the final time inside the timed region
        self.dev.project(np.ascontiguousarray(vals))
        self.dof_per_problem = w['cells'] * self.nd * self.nvar
        self.ndof_local = self.dof_per_problem * w['nprob']

    def run(self, n):
        if self.dd is not None:
            self.dd.run_steps(n, self.cfl, self.dx, self.Tf, self.Mdx2, graph=self.graph)
        elif self.graph:
            self.dev.run_steps_graphed(n, self.cfl, self.dx, self.Tf, self.Mdx2)
        else:
            self.dev.run_steps(n, self.cfl, self.dx, self.Tf, self.Mdx2)

    def prime(self):
        (self.dd if self.dd is not None else self.dev).max_speed()

    def sync_all(self):
        import torch.distributed as dist
        self.torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
            self.torch.cuda.synchronize()

    def _allmax(self, x):
        import torch.distributed as dist
        t = self.torch.tensor([x], dtype=self.torch.float64, device='cuda')
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def _allsum(self, t):
        import torch.distributed as dist
        t = t.to(self.torch.float64).sum()
        if self.world > 1:
            dist.all_reduce(t)
        return float(t.item())

    def timed(self, K, W, lib, local_rank):
        """W untimed warm-up steps, then exactly K steps between barriers + device synchronisation; CUDA events, max over
        ranks.  Returns the fields of one measurement."""
        torch = self.torch
        self.prime()
        self.run(W)
        self.sync_all()
        sampler = ClockSampler(local_rank) if self.rank == 0 else None
        it0 = self.dev.it.clone()
        l0 = lib.dg1d_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tw0 = time.time()
        e0.record()
        self.run(K)
        e1.record()
        self.sync_all()
        tw1 = time.time()
        launches = lib.dg1d_launch_count() - l0
        if self.graph:     # replayed kernels do not pass through the library's counter: 4 kernels per step (+ halo kernels)
            launches = K * (4 + (6 if (self.world > 1 and self.dd is not None) else 0))
        ms = self._allmax(e0.elapsed_time(e1))
        clocks = sampler.stop(tw0, tw1) if sampler else None
        state_ok = bool(torch.isfinite(self.dev.bufs[0]).all().item())
        # work actually done: steps executed per problem (an ensemble member that blows up -- the reference scheme has no
        # positivity fix, and about 1 % of the config-5 Riemann problems produce a negative pressure -- is passed through
        # afterwards and must not be counted)
        steps_done = self._allsum(self.dev.it - it0)
        value = steps_done * self.dof_per_problem * 3 / (ms * 1e-3)
        dead = int((self.dev.it - it0 < K).sum().item())
        if self.dd is not None:
            self.dd.check_transport()          # raises if a peer-memory wait timed out
        return dict(value=value, ms=ms, ms_per_step=ms / K, launches=int(launches), state_ok=state_ok, dead=dead, clocks=clocks)

    def e2e(self, K):
        """The same metric through the host-buffer API: pinned host state -> H2D -> pack, K steps with (t, dt) read back to
        the host after every step, unpack -> D2H; everything inside the timed region."""
        torch, dev = self.torch, self.dev
        host = torch.empty(dev.aos_shape, dtype=torch.float64).pin_memory()
        dev.download(host)
        ctl_host = torch.empty(2, dtype=torch.float64).pin_memory()
        self.sync_all()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        dev.upload(host)                                             # H2D of the inputs (pinned) into the staging area + pack
        dev.reset_control()
        self.prime()
        for _ in range(K):
            self.run(1)
            ctl_host[0:1].copy_(dev.t[0:1], non_blocking=True)      # the step's result read back every step
            ctl_host[1:2].copy_(dev.dt[0:1], non_blocking=True)
            torch.cuda.current_stream().synchronize()
        dev.download(host)                                           # unpack + D2H of the final state
        f1.record()
        self.sync_all()
        ems = self._allmax(f0.elapsed_time(f1))
        nbytes = self.ndof_local * 8
        steps_e = self._allsum(dev.it)            # reset_control() zeroed the counters above
        if self.dd is not None:
            self.dd.check_transport()
        return {'value': steps_e * self.dof_per_problem * 3 / (ems * 1e-3), 'unit': 'DOF-updates/s', 'h2d_bytes_per_step': nbytes / K,
                'd2h_bytes_per_step': nbytes / K + 16, 'ms_total': ems,
                'how': 'pinned host state -> H2D + pack once, K steps with (t, dt) read back after each, unpack + D2H once; bytes are '
                       'per GPU, state bytes amortised over the K steps; the two state copies cross PCIe at its own rate and cannot '
                       'overlap the steps (the first step needs the whole state for the CFL reduction, the last one ends before the copy back)'}

    def close(self):
        if self.dd is not None:
            self.dd.close()
        self.dd = self.dev = None
        self.torch.cuda.empty_cache()


def roofline_of(value, world, ms_per_step, hbm_peak, peak_src, name):
    """HBM roofline of the dominant kernel (the fused stage kernel: 3 launches per step).  64 B per DOF per step
    (stage 0: read u, write; stages 1,2: read u, read u^n, write) = 21.33 B per DOF-update."""
    per_gpu = value / world
    avg_launch_s = ms_per_step * 1e-3 / 3.0
    alg_bytes_per_launch = 64.0 * per_gpu * (ms_per_step * 1e-3) / 9.0       # DOF-updates per step / 3 = DOFs; 64 B each; 3 launches
    achieved = alg_bytes_per_launch / avg_launch_s / 1e9
    traffic, src = None, None
    tp = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(tp):
        tj = json.load(open(tp))
        traffic, src = tj.get(name), tj.get('_source')
    return {'bound': 'hbm', 'achieved': achieved, 'peak': hbm_peak, 'unit': 'GB/s', 'frac': achieved / hbm_peak, 'traffic': traffic,
            'traffic_source': (src if traffic is not None else None), 'peak_source': peak_src, 'kernel': 'dg1d::stage_kernel',
            'algorithmic_bytes_per_launch': alg_bytes_per_launch, 'avg_launch_us': avg_launch_s * 1e6,
            'frac_of_nominal_8000': achieved / 8000.0,
            'note': 'per GPU; launch time = timed region / (3K): the step-control kernel%s is inside it'
                    % (' and the halo exchange' if world > 1 else '')}


def reference_python_baseline():
    """The reference itself (dg1d/python compiled to bytecode under oracle/_ref/py) on one host core: BASELINE config 1."""
    from oracle import ref_python
    if not ref_python.available():
        return {'unavailable': 'oracle/_ref/py is absent (built by `make -C oracle refpy` where /root/reference exists)'}
    rate, secs, it, ok = ref_python.time_config1()
    return {'value': rate, 'unit': 'DOF-updates/s', 'cores': 1, 'kind': 'reference', 'seconds': secs, 'steps': it,
            'known_answers_ok': ok,
            'workload': 'config 1 in full: euler.py -ncell 200 -degree 1 -ic sod -Tf 0.2 (Rusanov, TVD limiter; the unmodified '
                        'reference, single-threaded Python/numpy; its rate does not depend on the mesh size, SURVEY F10)'}


def main():
    real_stdout = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=('ours', 'reference'))
    ap.add_argument('--workload', default='c4', choices=sorted(WORKLOADS))
    ap.add_argument('--degree', type=int, default=None)
    ap.add_argument('--cells', type=int, default=None, help='cells per GPU (override)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-secondary', action='store_true', help='skip the secondary workloads, the NCCL re-timing and the multi-GPU parity check')
    ap.add_argument('--post-shock', action='store_true',
                    help='c3: start from a state advanced past shock formation (t=0.25) at 2e4 cells and re-projected (SURVEY 8d)')
    ap.add_argument('--transport', choices=('auto', 'peer', 'nccl'), default='auto',
                    help='N>1: halo/CFL exchange through peer-memory mailboxes (NVLink stores) or NCCL; auto = peer if available')
    ap.add_argument('--python-dist', action='store_true', help='N>1: drive the exchanges from Python (torch.distributed) instead of the C loop')
    ap.add_argument('--graph', action='store_true', help='replay CUDA graphs of two steps instead of launching each kernel')
    a = ap.parse_args()

    def workload(name):
        w = dict(WORKLOADS[name])
        if name == a.workload:
            if a.degree is not None:
                w['degree'] = a.degree
            if a.cells is not None:
                w['cells'] = a.cells
        if name in ('c2', 'c2big') and w['degree'] == 4:
            w['cfl'] = 0.8      # the reference itself is unstable at 0.9 for degree 4
        return w

    w = workload(a.workload)
    rank = int(os.environ.get('RANK', '0')); world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    W = max(a.warmup, 3)
    K = a.steps
    hbm_peak, peak_src = peaks()
    base = {'metric': 'DG DOF-updates/sec per RK stage', 'unit': 'DOF-updates/s', 'n_gpus': a.gpus, 'steps': K, 'warmup': W,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': make_config(a.workload, w, world)}

    if a.impl == 'reference':
        if rank != 0:
            return
        # torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm is meant to use every host thread
        os.environ['OMP_NUM_THREADS'] = str(os.cpu_count() or 1)
        rate, info = cpu_port(w, steps=K, warmup=min(W, 3))
        out = dict(base)
        out.update({'impl': 'reference', 'value': rate, 'ms_per_step': info['seconds'] / info['steps'] * 1e3, 'n_gpus': a.gpus,
                    'gpu_launches': 0,
                    'cpu_baseline': {'value': rate, 'unit': 'DOF-updates/s', 'cores': info['cores'], 'kind': 'port', 'sample': info['sample']},
                    'e2e': {'value': rate, 'unit': 'DOF-updates/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
                    'note': 'the C restatement of the reference algorithm (oracle/, bit-checked against reference-generated goldens) '
                            'with all host threads on the sample named in config.cpu_reference_sample: the harder of the two CPU '
                            'baselines.  The reference itself (dg1d/python, one core, about 7e4 DOF-updates/s at any size) is timed '
                            'by the repo arm under cpu_baseline.reference_python'})
        _emit(real_stdout, out)
        return

    import torch
    import torch.distributed as dist
    from fem_b200 import _lib
    assert torch.cuda.is_available(), 'bench.py needs a GPU; there is no CPU fallback'
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    lib = _lib.load()
    extras = not a.no_secondary

    # ---- N > 1: multi-GPU parity first (tests/dist_check.py: split runs bit-compared with the serial oracle on rank 0,
    # both transports, including one-cell last tiles and the reference's own blow-up); a failure ends the run
    dist_parity = None
    if world > 1 and extras:
        sys.path.insert(0, os.path.join(ROOT, 'tests'))
        import dist_check
        msgs = []
        dist_parity = dist_check.run_all(('peer', 'nccl'), native=True, log=msgs.append)
        flag = torch.tensor([1 if (dist_parity is None or all(all(v.values()) for v in dist_parity.values())) else 0], device='cuda')
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            for m in msgs:
                print(m, file=sys.stderr)
        if int(flag.item()) == 0:
            if rank == 0:
                print('multi-GPU parity FAILED: %r' % (dist_parity,), file=sys.stderr)
            dist.destroy_process_group()
            sys.exit(3)

    # ---- the headline workload ------------------------------------------------------------------------------------------
    case = Case(a.workload, w, world, rank, transport=a.transport, native=not a.python_dist, post_shock=a.post_shock, graph=a.graph)
    if case.transport:
        base['transport'] = case.transport
    if case.state_note:
        base['state'] = case.state_note
    m = case.timed(K, W, lib, local_rank)
    e2e = None if a.no_e2e else case.e2e(K)
    case.close()

    # ---- N > 1: the same workload over NCCL (ncclSend/ncclRecv halo, ncclAllReduce) when the headline used peer memory ----
    nccl = None
    if world > 1 and extras and w['nprob'] == 1 and base.get('transport') != 'nccl':
        c2 = Case(a.workload, w, world, rank, transport='nccl')
        mn = c2.timed(K, W, lib, local_rank)
        c2.close()
        nccl = {'ms_per_step': mn['ms_per_step'], 'value': mn['value'], 'state_finite': mn['state_ok']}

    # ---- secondary workloads: short runs of the other configs on the same line ------------------------------------------------
    secondary = {}
    if extras:
        for name in SECONDARY:
            if name == a.workload:
                continue
            ws = workload(name)
            cs = Case(name, ws, world, rank, transport=a.transport)
            ks = 200 if ws['cells'] * ws['nprob'] <= 2_000_000 else 50      # the 8-microsecond kernels of c2 need more steps to time
            ms_ = cs.timed(ks, 5, lib, local_rank)
            cs.close()
            secondary[name] = {'workload': ws['desc'], 'value': ms_['value'], 'ms_per_step': ms_['ms_per_step'], 'steps': ks, 'warmup': 5,
                               'roofline_frac': roofline_of(ms_['value'], world, ms_['ms_per_step'], hbm_peak, peak_src, name)['frac'],
                               'state_finite': ms_['state_ok'], 'problems_stopped_early_rank0': ms_['dead'], 'clocks': ms_['clocks'],
                               'transport': cs.transport, 'l2_policy': make_config(name, ws, world)['l2_policy']}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    out = dict(base)
    out.update({'value': m['value'], 'ms_per_step': m['ms_per_step'], 'gpu_launches': m['launches'], 'state_finite': m['state_ok'],
                'problems_stopped_early_rank0': m['dead'],
                'roofline': roofline_of(m['value'], world, m['ms_per_step'], hbm_peak, peak_src, a.workload), 'clocks': m['clocks']})
    if e2e:
        out['e2e'] = e2e
    if nccl:
        out['ms_per_step_nccl'] = nccl['ms_per_step']
        out['nccl_transport'] = nccl
    if dist_parity is not None:
        out['dist_parity'] = dist_parity
    if secondary:
        out['workloads'] = secondary
    if not a.no_cpu_baseline and world == 1:
        os.environ.setdefault('OMP_NUM_THREADS', str(os.cpu_count() or 1))
        rate, info = cpu_port(w)
        out['cpu_baseline'] = {'value': rate, 'unit': 'DOF-updates/s', 'cores': info['cores'], 'kind': 'port', 'sample': info['sample'],
                               'reference_python': reference_python_baseline()}
    _emit(real_stdout, out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
