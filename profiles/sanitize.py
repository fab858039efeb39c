"""Small invocations of every kernel family (all Euler fluxes and limiters, every scalar PDE and boundary kind, the unfused
pieces, the norm): a quick all-kernels smoke, and the target for compute-sanitizer where that tool is available
(it was closed on the round-1 pool; out-of-bounds accesses are also hunted by tests/test_gpu_parity.py::test_no_write_outside_the_state_buffers)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200.device import Dg1dDevice

rng = np.random.default_rng(0)


def euler_state(nc, nd, nprob=1):
    u = np.zeros((3, nprob, nc, nd))
    x = np.linspace(0, 1, nc)
    rho = 1.0 + 0.3 * np.sin(6 * x); vel = 0.3 * np.cos(4 * x); pre = 1.0 + 0.2 * np.cos(3 * x)
    u[0, :, :, 0] = rho; u[1, :, :, 0] = rho * vel; u[2, :, :, 0] = pre / 0.4 + 0.5 * rho * vel ** 2
    for m in range(1, nd):
        u[:, :, :, m] = 0.01 / m * rng.standard_normal((3, nprob, nc))
    return u


for flux, k, char, nc, nprob in (('rusanov', 2, 1, 700, 1), ('rusanov', 1, 0, 300, 3), ('lfc', 2, 1, 260, 1), ('hllc', 1, 0, 130, 1),
                                 ('roe', 3, 1, 129, 1)):
    dev = Dg1dDevice('euler', flux, k, nc, limiter=True, char_lim=char, bc=('copy', 'copy'), nprob=nprob)
    dev.pack(euler_state(nc, k + 1, nprob))
    dev.max_speed()
    dev.run_steps(3, 0.1, 1.0 / nc, 1e9, 10.0 / nc ** 2)
    A, B = dev.new_state(), dev.new_state()
    dev.residual(dev.bufs[0], A); dev.limiter(dev.bufs[0], B, 0.0)
    dev.eval_points()
    assert torch.isfinite(dev.unpack()).all()
for pde, flux, k, lim, bc in (('linear', 'upwind', 1, False, ('periodic', 'periodic')), ('burger', 'roe', 2, True, ('periodic', 'periodic')),
                              ('varadv', 'upwind', 3, True, ('periodic', 'periodic')), ('burger', 'godunov', 2, True, ('inflow', 'outflow')),
                              ('linear', 'central', 4, False, ('outflow', 'inflow'))):
    nc = 515
    dev = Dg1dDevice(pde, flux, k, nc, limiter=lim, bc=bc)
    dev.set_mesh(0.0, 1.0 / nc)
    if 'inflow' in bc:
        dev.set_boundary([1.2, 0.9])
    u = 1.0 + 0.3 * rng.standard_normal((1, 1, nc, k + 1)) / (1 + np.arange(k + 1))
    dev.pack(u)
    dev.max_speed()
    dev.run_steps(3, 0.1, 1.0 / nc, 1e9, 5.0 / nc ** 2)
    V = np.ones((k + 3, k + 1)); w = np.ones(k + 3)
    dev.error_norm(np.zeros((1, nc, k + 3)), V, w, 1.0 / nc)
    dev.eval_points()
    assert torch.isfinite(dev.unpack()).all()
torch.cuda.synchronize()
print('sanitize.py: all kernels ran')
