"""oracle/ref_python.py runs the unmodified reference from its compiled form (oracle/_ref/py, built by `make -C oracle
refpy` where /root/reference exists; it travels to the GPU box).  When it is present it must reproduce the committed
goldens, which were recorded from the reference's sources: the CPU baseline bench.py times is then the real thing."""
import numpy as np
import pytest

from conftest import load_golden
from oracle import ref_python


@pytest.mark.skipif(not ref_python.available(), reason='oracle/_ref/py not built on this machine')
def test_compiled_reference_reproduces_goldens():
    g = load_golden('lin_sin2pi_k2_nc20')
    dg = ref_python.run_scalar(pde='linear', ncell=20, degree=2, cfl=0.9, Tf=0.1, ic='sin2pi', limit='no', num_flux='upwind')
    assert np.array_equal(dg.u1, g['u_final']) and np.array_equal(dg.u0, g['u0_final'])
    g = load_golden('eul_sod_k0_nc64')
    e = ref_python.run_euler(['-ncell', 64, '-degree', 0, '-ic', 'sod', '-Tf', 0.1, '-plot_freq', 100000])
    assert int(e['it']) == g['it']
    assert np.array_equal(np.array([e['rho1'], e['mom1'], e['ene1']]), g['u_final'])
