"""Runs the UNMODIFIED Python reference (cpraveen/fem dg1d/python) from its compiled form.

TEST INFRASTRUCTURE ONLY, like everything under oracle/: used by bench.py's cpu_baseline leg (and usable by tests).
oracle/Makefile (target refpy) compiles every file of dg1d/python/{scalar,euler} where it lies into CPython bytecode,
oracle/_ref/py/{scalar,euler}/<module>.pycode -- git-ignored, travels to the GPU box like the compiled C references.
A small importer below serves those files to `import`; the code objects run exactly as the sources do (same
interpreter version on both machines: same image).

matplotlib is not installed: a throw-away stub package is put on sys.path and input() is patched, as SURVEY.md
Appendix A and tests/golden/gen_golden.py do; no reference code is changed.
"""
import builtins
import importlib.abc
import importlib.util
import marshal
import os
import sys
import tempfile
import time
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
PY = os.path.join(_HERE, '_ref', 'py')
EULER_MODS = ['basis', 'limiter', 'gas', 'sod', 'lax', 'shuosher', 'riemann']
SCALAR_MODS = ['dgsolver', 'basis', 'limiter', 'linadv', 'burger', 'varadv', 'sin2pi', 'sin4pi', 'gauss', 'hat', 'mult']
_stub_dir = None


def available():
    return os.path.exists(os.path.join(PY, 'euler', 'euler.pycode')) and os.path.exists(os.path.join(PY, 'scalar', 'dgsolver.pycode'))


def _code(path):
    with open(path, 'rb') as f:
        return marshal.loads(f.read()[16:])       # 16-byte .pyc header: magic, flags, source mtime and size


class _Importer(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Serves <dir>/<name>.pycode to `import name` while it is on sys.meta_path."""

    def __init__(self, d):
        self.d = d

    def find_spec(self, name, path=None, target=None):
        if '.' not in name and os.path.exists(os.path.join(self.d, name + '.pycode')):
            return importlib.util.spec_from_loader(name, self)
        return None

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        exec(_code(os.path.join(self.d, module.__name__ + '.pycode')), module.__dict__)


def _stubs():
    global _stub_dir
    if _stub_dir is None:
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
        _stub_dir = d
    if _stub_dir not in sys.path:
        sys.path.insert(0, _stub_dir)
    builtins.input = lambda *a, **k: ''


def _enter(sub):
    _stubs()
    for n in EULER_MODS + SCALAR_MODS:
        sys.modules.pop(n, None)
    imp = _Importer(os.path.join(PY, sub))
    sys.meta_path.insert(0, imp)
    return imp


def run_euler(argv):
    """euler.py as a script (it is one: euler.py:15-28, 181-274) with these command-line flags; returns its globals
    (rho1, mom1, ene1, it, t, dx, ...).  -plot_freq large keeps the plotting stub out of the loop."""
    imp = _enter('euler')
    old = sys.argv
    sys.argv = ['euler.py'] + [str(a) for a in argv]
    g = {'__name__': '__main__', '__builtins__': builtins}
    try:
        exec(_code(os.path.join(imp.d, 'euler.pycode')), g)
        return g
    finally:
        sys.argv = old
        sys.meta_path.remove(imp)


def run_scalar(**kw):
    """dgsolver.DG(args).run() (dgsolver.py:281-294); returns the DG object after the run."""
    imp = _enter('scalar')
    try:
        import dgsolver
        a = dict(pde='linear', ncell=50, nrefine=0, degree=1, cfl=0.9, Tf=1.0, plot_freq=0, ic='sin2pi', limit='no', tvbM=0.0,
                 compute_error='no', num_flux='upwind')
        a.update(kw)
        dg = dgsolver.DG(types.SimpleNamespace(**a))
        dg.result = dg.run()
        return dg
    finally:
        sys.meta_path.remove(imp)


def time_config1():
    """BASELINE config 1, the reference's own CPU run, in full: `python euler.py -ncell 200 -degree 1 -ic sod -Tf 0.2`
    (Rusanov, TVD limiter, SSP-RK3; 292 steps).  Returns (DOF-updates/s per RK stage, seconds, steps, known-answer ok)."""
    t0 = time.perf_counter()
    g = run_euler(['-ncell', 200, '-degree', 1, '-ic', 'sod', '-Tf', 0.2, '-plot_freq', 100000])
    dt = time.perf_counter() - t0
    it = int(g['it'])
    ndof = 200 * 2 * 3
    # the survey's known answers for this run: it = 292, rho1[100] = [0.42674912866909137, -1.8062216555310572e-04]
    ok = it == 292 and abs(float(g['rho1'][100][0]) - 4.2674912866909137e-01) < 1e-13
    return ndof * 3 * it / dt, dt, it, ok
