import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


def golden_names(kind=None):
    names = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, '*.npz')))
    if kind == 'vectors':
        return [n for n in names if n.startswith('csibvec_')]
    if kind == 'evalpts':
        return [n for n in names if n.startswith('evalpts_')]
    names = [n for n in names if not n.startswith(('csibvec_', 'evalpts_'))]      # function-level vectors / plot data, not solver runs
    if kind == 'scalar':
        names = [n for n in names if n.startswith(('lin_', 'bur_', 'var_'))]
    elif kind == 'euler':
        names = [n for n in names if n.startswith(('eul_', 'csib_', 'dg2d_'))]
    return names


def load_golden(name):
    d = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    out = {}
    for k in d.files:
        v = d[k]
        out[k] = v.item() if v.ndim == 0 else v
    return out


def rel_linf(a, b):
    """Relative L-infinity error, the norm north_star states (1e-12 per stage, 1e-10 final)."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    den = np.max(np.abs(b))
    if den == 0.0:
        den = 1.0
    return float(np.max(np.abs(a - b)) / den)
