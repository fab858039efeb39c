#!/usr/bin/env python3
"""Generate golden vectors by executing the UNMODIFIED reference (cpraveen/fem dg1d/python).

Container-only tool: it reads /root/reference, which does not exist on the GPU box.  Nothing in
tests/, bench.py or the product imports this file; the tests only read the .npz files it writes.

How the reference is driven:
  * matplotlib is absent, so a throw-away stub package is put first on sys.path and
    builtins.input is patched; no reference file is modified or copied.
  * scalar: `dgsolver.DG(args)` is instantiated and its own `solve()` runs; the instance's
    `compute_residual` / `apply_limiter` are wrapped with recorders, so every snapshot is
    taken from the reference's own arrays while its own driver loop runs
    (dg1d/python/scalar/dgsolver.py:255-278).
  * euler: `euler.py` is a module-level script; it is run with runpy and a sys.settrace line
    tracer copies the global state arrays at euler.py:235 (residual complete, before the RK
    update), :239 (after the update, before the limiter) and at the next line event after the
    limiter loop (:190 or :274)  (dg1d/python/euler/euler.py:181-274).

Usage:  python tests/golden/gen_golden.py [case-name-substring ...]
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
REPO = os.path.dirname(os.path.dirname(OUT))


def _install_stubs():
    d = tempfile.mkdtemp(prefix='mplstub_')
    os.makedirs(os.path.join(d, 'matplotlib'))
    with open(os.path.join(d, 'matplotlib', '__init__.py'), 'w') as f:
        f.write('from unittest.mock import MagicMock\nrcParams = {}\n')
    with open(os.path.join(d, 'matplotlib', 'pyplot.py'), 'w') as f:
        f.write('from unittest.mock import MagicMock\n_m = MagicMock()\n'
                'def subplots(nrows=1, ncols=1, **kw):\n'
                '    ax = [MagicMock() for _ in range(ncols)]\n'
                '    for a in ax: a.plot.return_value = (MagicMock(),)\n'
                '    return MagicMock(), ax\n'
                'def __getattr__(name): return getattr(_m, name)\n')
    sys.path.insert(0, d)
    builtins.input = lambda *a, **k: ''


def _purge(names):
    for n in names:
        sys.modules.pop(n, None)


SCALAR_MODS = ['dgsolver', 'basis', 'limiter', 'linadv', 'burger', 'varadv', 'sin2pi', 'sin4pi',
               'gauss', 'hat', 'mult']
EULER_MODS = ['basis', 'limiter', 'gas', 'sod', 'lax', 'shuosher', 'riemann']


# --------------------------------------------------------------------------------------------
# scalar
# --------------------------------------------------------------------------------------------
def run_scalar(name, pde='linear', ic='sin2pi', degree=1, ncell=20, cfl=0.9, Tf=0.1, limit='no',
               tvbM=0.0, num_flux='upwind', record_steps=(0, 1, 2), compute_error='yes'):
    """record_steps: time-step indices whose three stages are stored in full."""
    _purge(SCALAR_MODS + EULER_MODS)
    sys.path.insert(0, os.path.join(REF, 'scalar'))
    try:
        import dgsolver
        args = types.SimpleNamespace(pde=pde, ncell=ncell, nrefine=0, degree=degree, cfl=cfl, Tf=Tf,
                                     plot_freq=0, ic=ic, limit=limit, tvbM=tvbM,
                                     compute_error=compute_error, num_flux=num_flux)
        dg = dgsolver.DG(args)
        rec = {'stage_in': [], 'res': [], 'pre_lim': [], 'post_lim': [], 'stage_id': []}
        state = {'n': 0}
        orig_res, orig_lim = dg.compute_residual, dg.apply_limiter
        want = set(record_steps)

        def compute_residual():
            step = state['n'] // 3
            if step in want:
                rec['stage_in'].append(dg.u1.copy())
            orig_res()
            if step in want:
                rec['res'].append(dg.res.copy())
                rec['stage_id'].append(state['n'])
            state['n'] += 1

        def apply_limiter():
            step = (state['n'] - 1) // 3
            if step in want:
                rec['pre_lim'].append(dg.u1.copy())
            orig_lim()
            if step in want:
                rec['post_lim'].append(dg.u1.copy())

        dg.compute_residual, dg.apply_limiter = compute_residual, apply_limiter
        dg.setup()
        dg.set_ic()
        u_init = dg.u1.copy()
        dg.solve()
        out = dict(kind='scalar', pde=pde, ic=ic, degree=degree, ncell=ncell, cfl=cfl, Tf=Tf, limit=limit,
                   tvbM=tvbM, num_flux=num_flux, u_init=u_init, u_final=dg.u1.copy(),
                   u0_final=dg.u0.copy(), nstages=state['n'],
                   Vf=dg.Vf, Vg=dg.Vg, Vu=dg.Vu, bm=dg.bm, bp=dg.bp, xg=dg.xg, wg=dg.wg,
                   ark=dgsolver.ark, brk=dgsolver.brk, dx=dg.dx, Mdx2=dg.Mdx2, cfl_eff=dg.cfl,
                   xmin=dg.xmin, xmax=dg.xmax,
                   stage_id=np.array(rec['stage_id']), stage_in=np.array(rec['stage_in']),
                   res=np.array(rec['res']), pre_lim=np.array(rec['pre_lim']),
                   post_lim=np.array(rec['post_lim']))
        if compute_error == 'yes':
            nc, dx, err = dg.compute_error_norm()
            out['err'] = err
        np.savez_compressed(os.path.join(OUT, name + '.npz'), **out)
        print('%-34s stages=%5d  u_final[0]=%s' % (name, state['n'], dg.u1[0]))
    finally:
        sys.path.remove(os.path.join(REF, 'scalar'))


# --------------------------------------------------------------------------------------------
# euler
# --------------------------------------------------------------------------------------------
def run_euler(name, ic='sod', degree=1, ncell=100, cfl=0.9, Tf=0.0, char_lim=0, tvbM=0.0,
              record_steps=(0, 1, 2), riemann=None, csib_flux=None):
    """riemann=(vl, vr, xs): replace sod's states through the reference's own riemann_data
    (dg1d/python/euler/riemann.py:4-13), as SURVEY.md 8(c) prescribes for ensemble members.
    csib_flux in {lf, ecusp, hllc, ausmdv, lfc}: the reference loop with gas.numflux replaced by the C
    sibling's own flux function (dg1d/c/euler/src/flux.c compiled unmodified into oracle/_ref),
    the same way SURVEY.md 8(c) plugs a flux into the loop."""
    _purge(SCALAR_MODS + EULER_MODS)
    edir = os.path.join(REF, 'euler')
    sys.path.insert(0, edir)
    script = os.path.join(edir, 'euler.py')
    try:
        if csib_flux is not None:
            import gas                       # first import after the purge: sod/lax/shuosher re-export it
            sys.path.insert(0, REPO)
            from oracle import dg1d_oracle as _orc
            if csib_flux == 'roe':    # PDE::roe_flux of dg2d/models/euler/pde.h:209-285 compiled unmodified (oracle/ref_dg2d_roe.cc)
                gas.numflux = lambda ul, ur: _orc.roe_ref_flux([float(v) for v in ul], [float(v) for v in ur])
            else:
                gas.numflux = lambda ul, ur: _orc.csib_ref_flux(csib_flux, [float(v) for v in ul], [float(v) for v in ur])
        if riemann is not None:
            import sod
            import riemann as riem
            vl, vr, xs = riemann
            sod.initial_condition = lambda x: riem.riemann_data(list(vl), list(vr), xs, x)
        argv = ['euler.py', '-ncell', str(ncell), '-degree', str(degree), '-ic', ic, '-cfl', repr(cfl),
                '-Tf', repr(Tf), '-char_lim', str(char_lim), '-tvbM', repr(tvbM),
                '-plot_freq', '100000000']
        rec = {'stage_in': [], 'res': [], 'pre_lim': [], 'post_lim': [], 'lam': [], 'dt': [],
               'stage_id': [], 'u_init': None}
        st = {'n': 0, 'pending_post': False}
        want = set(record_steps)

        def grab(g):
            return np.array([g['rho1'], g['mom1'], g['ene1']])

        def tracer(frame, event, arg):
            if frame.f_code.co_filename != script:
                return None
            return local

        def local(frame, event, arg):
            if event != 'line':
                return local
            g = frame.f_globals
            ln = frame.f_lineno
            if st['pending_post'] and ln in (190, 274):
                st['pending_post'] = False
                step = (st['n'] - 1) // 3
                if step in want:
                    rec['post_lim'].append(grab(g))
            if ln == 172 and rec['u_init'] is None:
                rec['u_init'] = grab(g)
            elif ln == 187:
                rec['dt'].append(g['dt'])
                rec['lam'].append(g['lam'])
            elif ln == 235:
                step = st['n'] // 3
                if step in want:
                    rec['stage_in'].append(grab(g))
                    rec['res'].append(np.array([g['resr'], g['resm'], g['rese']]))
                    rec['stage_id'].append(st['n'])
            elif ln == 239:
                step = st['n'] // 3
                if step in want:
                    rec['pre_lim'].append(grab(g))
                st['n'] += 1
                st['pending_post'] = True
            return local

        old_argv = sys.argv
        sys.argv = argv
        sys.settrace(tracer)
        try:
            g = runpy.run_path(script, run_name='__main__')
        finally:
            sys.settrace(None)
            sys.argv = old_argv
        out = dict(kind='euler', ic=ic, degree=degree, ncell=ncell, cfl=cfl, Tf=g['Tf'], char_lim=char_lim,
                   tvbM=tvbM, u_init=rec['u_init'], u_final=grab(g),
                   it=g['it'], t=g['t'], nstages=st['n'],
                   Vf=g['Vf'], Vg=g['Vg'], Vu=g['Vu'], bm=g['bm'], bp=g['bp'], xg=g['xg'], wg=g['wg'],
                   ark=g['ark'], brk=g['brk'], sqrt3=g['sqrt3'], isqrt3=g['isqrt3'],
                   dx=g['dx'], Mdx2=g['Mdx2'], cfl_eff=g['cfl'], xmin=g['xmin'], xmax=g['xmax'],
                   dt=np.array(rec['dt']), lam=np.array(rec['lam']),
                   stage_id=np.array(rec['stage_id']), stage_in=np.array(rec['stage_in']),
                   res=np.array(rec['res']), pre_lim=np.array(rec['pre_lim']),
                   post_lim=np.array(rec['post_lim']))
        if csib_flux is not None:
            out['num_flux'] = csib_flux
        if riemann is not None:
            out['riemann_vl'] = np.array(riemann[0])
            out['riemann_vr'] = np.array(riemann[1])
            out['riemann_xs'] = riemann[2]
        np.savez_compressed(os.path.join(OUT, name + '.npz'), **out)
        print('%-34s it=%5d t=%.17g  rho1[nc//2]=%s' % (name, g['it'], g['t'], g['rho1'][ncell // 2]))
    finally:
        sys.path.remove(edir)
        _purge(EULER_MODS)


def csib_flux_vectors(name, n=4000, seed=20261107):
    """Input/output vectors of the C sibling's flux functions themselves (flux.c, compiled into oracle/_ref):
    random sub/supersonic pairs, near-equal pairs, equal pairs, strong jumps, every HLLC/AUSMDV/ECUSP branch."""
    sys.path.insert(0, REPO)
    from oracle import dg1d_oracle as _orc
    rng = np.random.default_rng(seed)

    def state():
        d = rng.uniform(0.05, 4.0); p = rng.uniform(0.02, 4.0); c = np.sqrt(1.4 * p / d)
        v = rng.choice([0.0, 1.0, 1.0, 1.0]) * rng.uniform(-3.0, 3.0) * c
        return np.array([d, d * v, p / (1.4 - 1.0) + 0.5 * d * v * v])

    ul = np.empty((n, 3)); ur = np.empty((n, 3))
    for i in range(n):
        a = state()
        m = i % 5
        b = state() if m < 2 else a * (1.0 + 1e-3 * rng.standard_normal(3)) if m == 2 else a.copy() if m == 3 \
            else a * np.array([rng.uniform(0.01, 100.0), rng.uniform(-2.0, 2.0), rng.uniform(0.5, 50.0)])
        if _orc.euler_numflux('rusanov', b, b)[1] != _orc.euler_numflux('rusanov', b, b)[1]:   # negative pressure
            b = state()
        ul[i], ur[i] = a, b
    out = dict(kind='csib_flux', ul=ul, ur=ur)
    for f in _orc.CSIB_FLUXES:
        out[f] = np.array([_orc.csib_ref_flux(f, ul[i], ur[i]) for i in range(n)])
    out['roe'] = np.array([_orc.roe_ref_flux(ul[i], ur[i]) for i in range(n)])      # dg2d/models/euler/pde.h:209-285
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **out)
    print('%-34s %d pairs, finite outputs: %s' % (name, n, {f: int(np.isfinite(out[f]).all(1).sum()) for f in _orc.CSIB_FLUXES}))


def ensemble_states(n, seed=20260101):
    """SURVEY.md 8(d) C5 state distribution."""
    rng = np.random.default_rng(seed)
    rl = rng.uniform(0.5, 2.0, n); pl = rng.uniform(0.5, 2.0, n)
    rr = rng.uniform(0.1, 0.5, n); pr = rng.uniform(0.05, 0.5, n)
    ul = rng.uniform(-0.2, 0.2, n); ur = rng.uniform(-0.2, 0.2, n)
    return np.stack([rl, ul, pl], 1), np.stack([rr, ur, pr], 1)


CASES = []


def case(fn, name, **kw):
    CASES.append((fn, name, kw))


# ---- scalar: config 2 (linear advection, degrees 1-4), plus every flux / IC / limiter branch
for k, c in ((1, 0.9), (2, 0.9), (3, 0.9), (4, 0.8)):
    case(run_scalar, 'lin_sin2pi_k%d_nc20' % k, degree=k, ncell=20, cfl=c, Tf=0.1)
case(run_scalar, 'lin_sin2pi_k3_nc40_T1', degree=3, ncell=40, Tf=1.0, record_steps=(0,))
case(run_scalar, 'lin_sin2pi_k0_nc16', degree=0, ncell=16, Tf=0.2)
case(run_scalar, 'lin_central_k2_nc16', degree=2, ncell=16, Tf=0.05, num_flux='central')
case(run_scalar, 'lin_hat_k1_nc40_tvd', ic='hat', degree=1, ncell=40, Tf=0.2, limit='yes')
case(run_scalar, 'lin_hat_k2_nc40_tvb', ic='hat', degree=2, ncell=40, Tf=0.2, limit='yes', tvbM=20.0)
case(run_scalar, 'lin_mult_k3_nc100_tvb', ic='mult', degree=3, ncell=100, Tf=0.2, limit='yes', tvbM=50.0,
     record_steps=(0, 1, 40))
case(run_scalar, 'lin_gauss_k2_nc30', ic='gauss', degree=2, ncell=30, Tf=0.1)
case(run_scalar, 'lin_sin4pi_k1_nc30_tvd', ic='sin4pi', degree=1, ncell=30, Tf=0.1, limit='yes')
# ---- scalar: variable-coefficient advection (varadv.py)
case(run_scalar, 'var_sin4pi_k1_nc40', pde='varadv', ic='sin4pi', degree=1, ncell=40, Tf=0.2)
case(run_scalar, 'var_mult_k2_nc80_tvb', pde='varadv', ic='mult', degree=2, ncell=80, Tf=0.3, limit='yes', tvbM=20.0,
     record_steps=(0, 1, 30))
case(run_scalar, 'var_sin2pi_k3_nc24', pde='varadv', ic='sin2pi', degree=3, ncell=24, Tf=0.1)
# ---- scalar: config 3 (Burgers + TVB, degree 2, shock formation at t = 1/(2 pi))
for M in (0.0, 10.0):
    case(run_scalar, 'bur_roe_k2_nc100_M%d' % M, pde='burger', degree=2, ncell=100, Tf=0.3, limit='yes',
         tvbM=M, num_flux='roe', record_steps=(0, 1, 200, 201, 300))
# config 3 at the largest size the reference finishes in minutes (parity at nc = 100 ... 1000)
case(run_scalar, 'bur_roe_k2_nc1000_M10', pde='burger', degree=2, ncell=1000, Tf=0.3, limit='yes', tvbM=10.0, num_flux='roe',
     record_steps=(0, 2000))
case(run_scalar, 'bur_god_k1_nc50_tvd', pde='burger', degree=1, ncell=50, Tf=0.25, limit='yes',
     num_flux='godunov', record_steps=(0, 1, 60))
case(run_scalar, 'bur_upw_k2_nc40_nolim', pde='burger', degree=2, ncell=40, Tf=0.1, num_flux='upwind')
case(run_scalar, 'bur_cen_k1_nc40_nolim', pde='burger', degree=1, ncell=40, Tf=0.05, num_flux='central')
case(run_scalar, 'bur_god_k3_nc64_tvb', pde='burger', degree=3, ncell=64, Tf=0.3, limit='yes', tvbM=5.0,
     num_flux='godunov', record_steps=(0, 150))
# ---- euler: config 1 (the reference CPU run) and its TVB / characteristic variants
case(run_euler, 'eul_sod_k1_nc200', ic='sod', degree=1, ncell=200, Tf=0.2, record_steps=(0, 1, 2, 100, 291))
case(run_euler, 'eul_sod_k1_nc200_M100', ic='sod', degree=1, ncell=200, Tf=0.2, tvbM=100.0,
     record_steps=(0, 1, 100))
case(run_euler, 'eul_sod_k1_nc100_char', ic='sod', degree=1, ncell=100, Tf=0.2, char_lim=1,
     record_steps=(0, 1, 50))
case(run_euler, 'eul_sod_k2_nc100_char_M100', ic='sod', degree=2, ncell=100, Tf=0.2, char_lim=1, tvbM=100.0,
     record_steps=(0, 1, 50, 120))
case(run_euler, 'eul_sod_k0_nc64', ic='sod', degree=0, ncell=64, Tf=0.1)
case(run_euler, 'eul_sod_k3_nc50_char', ic='sod', degree=3, ncell=50, Tf=0.1, char_lim=1, tvbM=10.0,
     record_steps=(0, 1, 40))
case(run_euler, 'eul_sod_k4_nc40_char_M10', ic='sod', degree=4, ncell=40, Tf=0.1, char_lim=1, tvbM=10.0, record_steps=(0, 1, 60))
case(run_euler, 'eul_lax_k4_nc32_cons', ic='lax', degree=4, ncell=32, Tf=0.08, char_lim=0, tvbM=0.0, record_steps=(0, 40))
case(run_euler, 'eul_lax_k1_nc100', ic='lax', degree=1, ncell=100, Tf=0.3, record_steps=(0, 1, 30))
case(run_euler, 'eul_lax_k2_nc100_char', ic='lax', degree=2, ncell=100, Tf=0.3, char_lim=1, tvbM=50.0,
     record_steps=(0, 1, 30))
# ---- euler: config 4 (Shu-Osher, degree 2, characteristic TVB limiter, README parameters)
case(run_euler, 'eul_shu_k2_nc100_char_M100', ic='shuosher', degree=2, ncell=100, Tf=0.2, char_lim=1,
     tvbM=100.0, record_steps=(0, 1, 2, 52))
# nc=200 to the module's Tf=1.8: the reference itself blows up (negative pressure -> NaN) at it=624;
# kept as the NaN-propagation edge case.  The finite long run stops at t=0.9.
case(run_euler, 'eul_shu_k2_nc200_char_M100_T18', ic='shuosher', degree=2, ncell=200, Tf=1.8, char_lim=1,
     tvbM=100.0, record_steps=(0, 500))
case(run_euler, 'eul_shu_k2_nc200_char_M100_T09', ic='shuosher', degree=2, ncell=200, Tf=0.9, char_lim=1,
     tvbM=100.0, record_steps=(0, 300))
# config 4 at the README's mesh (README.md:40-41: 400 cells), stopped at t = 0.4
case(run_euler, 'eul_shu_k2_nc400_char_M100_T04', ic='shuosher', degree=2, ncell=400, Tf=0.4, char_lim=1, tvbM=100.0,
     record_steps=(0, 300))
case(run_euler, 'eul_shu_k2_nc100_cons_M100', ic='shuosher', degree=2, ncell=100, Tf=0.2, char_lim=0,
     tvbM=100.0, record_steps=(0, 52))
# ---- euler: config 5 (randomised Riemann ensemble members; 8 sampled problems at nc=256)
_vl, _vr = ensemble_states(4096)
for j in (0, 1, 17, 511, 512, 2048, 4000, 4095):
    case(run_euler, 'eul_ens%04d_k1_nc256' % j, ic='sod', degree=1, ncell=256, Tf=0.2,
         riemann=(tuple(_vl[j]), tuple(_vr[j]), 0.5), record_steps=(0, 1))
# ---- euler: the C sibling's flux family in the reference loop; needs oracle/_ref (make -C oracle)
for _f in ('lf', 'ecusp', 'hllc', 'ausmdv', 'lfc'):
    case(run_euler, 'csib_%s_sod_k1_nc100' % _f, ic='sod', degree=1, ncell=100, Tf=0.1, csib_flux=_f,
         record_steps=(0, 1, 40))
    case(run_euler, 'csib_%s_shu_k2_nc100_char_M100' % _f, ic='shuosher', degree=2, ncell=100, Tf=0.1, char_lim=1,
         tvbM=100.0, csib_flux=_f, record_steps=(0, 20))
case(run_euler, 'csib_lfc_lax_k3_nc60_char', ic='lax', degree=3, ncell=60, Tf=0.1, char_lim=1, tvbM=10.0, csib_flux='lfc',
     record_steps=(0, 30))
case(run_euler, 'csib_ausmdv_lax_k2_nc100', ic='lax', degree=2, ncell=100, Tf=0.2, csib_flux='ausmdv', record_steps=(0, 60))
case(csib_flux_vectors, 'csibvec_flux_pairs')
# ---- euler: the reference's own Roe flux (dg2d/models/euler/pde.h compiled unmodified) in the reference loop
case(run_euler, 'dg2d_roe_sod_k1_nc100', ic='sod', degree=1, ncell=100, Tf=0.1, csib_flux='roe', record_steps=(0, 1, 40))
case(run_euler, 'dg2d_roe_shu_k2_nc100_char_M100', ic='shuosher', degree=2, ncell=100, Tf=0.1, char_lim=1, tvbM=100.0,
     csib_flux='roe', record_steps=(0, 20))
case(run_euler, 'dg2d_roe_lax_k3_nc60_char', ic='lax', degree=3, ncell=60, Tf=0.1, char_lim=1, tvbM=10.0, csib_flux='roe',
     record_steps=(0, 30))

if __name__ == '__main__':
    if not os.path.isdir(REF):
        sys.exit('reference tree %s not present: goldens can only be generated in the build container' % REF)
    _install_stubs()
    sel = sys.argv[1:]
    for fn, name, kw in CASES:
        if sel and not any(s in name for s in sel):
            continue
        fn(name, **kw)
