#!/usr/bin/env python3
"""Golden vectors for the uniform-point evaluation (second half): what the UNMODIFIED reference hands to its
plotting calls -- `f = Vu @ u1[i,:]` per cell and the running min/max that become the axis limits
(dg1d/python/scalar/dgsolver.py:125-167), and for Euler density, velocity m/r and pressure (dg1d/python/euler/euler.py:77-148).

Container-only tool (reads /root/reference).  The reference runs with plotting ON; matplotlib is replaced by a stub that
RECORDS the arrays given to ax.plot / line.set_ydata and the limits given to plt.axis / ax.set_ylim.  `-plot_freq` is set
beyond the number of steps, so the reference draws twice: the initial condition (init_plot) and the final time
(update_plot fires when |Tf - t| < 1e-13).  No reference file is modified or copied.

Usage: python tests/golden/gen_eval_golden.py
"""
import builtins
import os
import runpy
import sys
import tempfile
import types

import numpy as np

REF = '/root/reference/dg1d/python'
OUT = os.path.dirname(os.path.abspath(__file__))

STUB = '''
import numpy as np
REC = {'plot': [], 'ydata': [], 'axis': [], 'ylim': []}
class Line:
    def __init__(self, ax): self.ax = ax
    def set_ydata(self, f): REC['ydata'].append((self.ax, np.array(f, dtype=np.float64).copy()))
class Manager:
    def set_window_title(self, *a): pass
class Canvas:
    manager = Manager()
class Ax:
    def __init__(self, i): self.i = i
    def plot(self, x, f, **kw):
        REC['plot'].append((self.i, np.array(x, dtype=np.float64).copy(), np.array(f, dtype=np.float64).copy()))
        return (Line(self.i),)
    def set_ylim(self, lim): REC['ylim'].append((self.i, [float(v) for v in lim]))
    def __getattr__(self, name): return lambda *a, **k: None
class Fig:
    canvas = Canvas()
    def add_subplot(self, *a): return Ax(0)
def figure(*a, **k): return Fig()
def subplots(nrows=1, ncols=1, **kw): return Fig(), [Ax(i) for i in range(ncols)]
def axis(lim): REC['axis'].append([float(v) for v in lim])
def __getattr__(name): return lambda *a, **k: None
'''


def install_stub():
    d = tempfile.mkdtemp(prefix='mplrec_')
    os.makedirs(os.path.join(d, 'matplotlib'))
    open(os.path.join(d, 'matplotlib', '__init__.py'), 'w').write('rcParams = {}\n')
    open(os.path.join(d, 'matplotlib', 'pyplot.py'), 'w').write(STUB)
    sys.path.insert(0, d)
    builtins.input = lambda *a, **k: ''
    import matplotlib.pyplot as plt
    return plt.REC


def purge():
    for n in ['dgsolver', 'basis', 'limiter', 'linadv', 'burger', 'varadv', 'sin2pi', 'sin4pi', 'gauss', 'hat', 'mult', 'gas', 'sod',
              'lax', 'shuosher', 'riemann']:
        sys.modules.pop(n, None)


def clear(rec):
    for v in rec.values():
        del v[:]


def run_scalar(rec, name, **kw):
    purge(); clear(rec)
    d = os.path.join(REF, 'scalar')
    sys.path.insert(0, d)
    try:
        import dgsolver
        a = dict(pde='linear', ncell=20, nrefine=0, degree=1, cfl=0.9, Tf=0.1, plot_freq=10 ** 9, ic='sin2pi', limit='no', tvbM=0.0,
                 compute_error='no', num_flux='upwind')
        a.update(kw)
        dg = dgsolver.DG(types.SimpleNamespace(**a))
        dg.setup(); dg.set_ic()
        u_init = dg.u1.copy()
        dg.u1[:, :] = 0.0                       # run() projects the IC again
        dg.run()
        nc = dg.nc
        assert len(rec['plot']) == nc and len(rec['ydata']) == nc and len(rec['axis']) == 2, (len(rec['plot']), len(rec['ydata']), rec['axis'])
        out = dict(kind='scalar', u_init=u_init, u_final=dg.u1.copy(), Vu=dg.Vu, xu=dg.xu, dx=dg.dx, xmin=dg.xmin,
                   x=np.array([p[1] for p in rec['plot']]), f_init=np.array([p[2] for p in rec['plot']]),
                   f_final=np.array([y[1] for y in rec['ydata']]), axis_init=np.array(rec['axis'][0]), axis_final=np.array(rec['axis'][1]),
                   **{k: v for k, v in a.items()})
        np.savez_compressed(os.path.join(OUT, name + '.npz'), **out)
        print('%-30s nu=%d f_final[0]=%s axis=%s' % (name, dg.Vu.shape[0], out['f_final'][0], rec['axis'][1]))
    finally:
        sys.path.remove(d)


def run_euler(rec, name, ic='sod', degree=1, ncell=100, cfl=0.9, Tf=0.0, char_lim=0, tvbM=0.0):
    purge(); clear(rec)
    d = os.path.join(REF, 'euler')
    sys.path.insert(0, d)
    old = sys.argv
    sys.argv = ['euler.py', '-ncell', str(ncell), '-degree', str(degree), '-ic', ic, '-cfl', repr(cfl), '-Tf', repr(Tf),
                '-char_lim', str(char_lim), '-tvbM', repr(tvbM), '-plot_freq', '1000000000']
    try:
        g = runpy.run_path(os.path.join(d, 'euler.py'), run_name='__main__')
        nc = ncell
        assert len(rec['plot']) == 3 * nc and len(rec['ydata']) == 3 * nc and len(rec['ylim']) == 6
        pl = {i: np.array([p[2] for p in rec['plot'] if p[0] == i]) for i in range(3)}
        yd = {i: np.array([y[1] for y in rec['ydata'] if y[0] == i]) for i in range(3)}
        out = dict(kind='euler', ic=ic, degree=degree, ncell=ncell, cfl=cfl, Tf=Tf, char_lim=char_lim, tvbM=tvbM,
                   u_final=np.array([g['rho1'], g['mom1'], g['ene1']]), Vu=g['Vu'], xu=g['xu'], dx=g['dx'], xmin=g['xmin'],
                   x=np.array([p[1] for p in rec['plot'] if p[0] == 0]),
                   r_init=pl[0], v_init=pl[1], p_init=pl[2], r_final=yd[0], v_final=yd[1], p_final=yd[2],
                   ylim_init=np.array([l[1] for l in rec['ylim'][:3]]), ylim_final=np.array([l[1] for l in rec['ylim'][3:]]),
                   it=g['it'], t=g['t'])
        np.savez_compressed(os.path.join(OUT, name + '.npz'), **out)
        print('%-30s nu=%d it=%d r_final[0]=%s ylim=%s' % (name, g['Vu'].shape[0], g['it'], yd[0][0], rec['ylim'][3:]))
    finally:
        sys.argv = old
        sys.path.remove(d)


def main():
    rec = install_stub()
    run_scalar(rec, 'evalpts_lin_k0_nc16', degree=0, ncell=16, Tf=0.2)
    run_scalar(rec, 'evalpts_lin_k1_nc20', degree=1, ncell=20, Tf=0.1)
    run_scalar(rec, 'evalpts_lin_k2_nc20', degree=2, ncell=20, Tf=0.1)
    run_scalar(rec, 'evalpts_lin_k3_nc20', degree=3, ncell=20, Tf=0.1)
    run_scalar(rec, 'evalpts_lin_k4_nc20', degree=4, ncell=20, Tf=0.1, cfl=0.8)
    run_scalar(rec, 'evalpts_bur_k2_nc100_M10', pde='burger', degree=2, ncell=100, Tf=0.3, limit='yes', tvbM=10.0, num_flux='roe')
    run_scalar(rec, 'evalpts_lin_hat_k1_nc40', degree=1, ncell=40, Tf=0.2, ic='hat', limit='yes')
    run_euler(rec, 'evalpts_eul_sod_k1_nc200', ic='sod', degree=1, ncell=200, Tf=0.2)
    run_euler(rec, 'evalpts_eul_shu_k2_nc100', ic='shuosher', degree=2, ncell=100, Tf=0.2, char_lim=1, tvbM=100.0)
    run_euler(rec, 'evalpts_eul_lax_k3_nc50', ic='lax', degree=3, ncell=50, Tf=0.1, char_lim=1, tvbM=10.0)
    run_euler(rec, 'evalpts_eul_sod_k0_nc64', ic='sod', degree=0, ncell=64, Tf=0.1)


if __name__ == '__main__':
    main()
