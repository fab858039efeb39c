#!/usr/bin/env python3
"""Config 1 (the reference's CPU run: Sod, degree 1, 200 cells, Rusanov, TVD, to t=0.2; 292 steps) end to end:
IC projection + time loop + download, eager launches vs CUDA-graph replay.  The reference takes 13.6 s."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200.euler import Euler, make_parser

for graph in (False, True, False, True):
    e = Euler(make_parser().parse_args(['-ncell', '200', '-degree', '1', '-ic', 'sod', '-Tf', '0.2']))
    e.setup(); e.set_ic(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    e.dev.reset_control()
    it, t = e.dev.solve(e.cfl, e.dx, e.Tf, e.Mdx2, graph=graph)
    u = e.u1
    dt = time.perf_counter() - t0
    print('graph=%s: it=%d t=%.3f  %.2f ms wall  (%.1f us/step)  rho1[100]=%r' % (graph, it[0], t[0], dt * 1e3, dt / it[0] * 1e6, u[0, 100, 0]))
