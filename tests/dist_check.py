#!/usr/bin/env python3
"""Multi-GPU parity check: every rank owns a block of the mesh (halo exchange per stage, max-reduction of the wave speed
per step); the gathered result must be bit-identical to the CPU oracle's serial run.

Run by hand under torchrun (not collected by pytest: it needs N GPUs), or called by bench.py at N > 1 (`run_all`):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dist_check.py

Cases: Euler chain (Shu-Osher, degree 2, characteristic limiter, copy ends), Burgers ring (degree 2, TVB, periodic), the
same two with a block size of 2*TILE+1 cells per rank (the last tile of a block then owns ONE cell, so the cell the halo
sends belongs to the tile before it), and the reference's own blow-up (Shu-Osher on 200 cells run to t = 1.8: a NaN wave
speed ends the loop, and must end it at the same step on every rank and with either transport)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import ic as ic_mod  # noqa: E402
from fem_b200.distributed import DistributedDg1d, block_partition  # noqa: E402
from oracle import dg1d_oracle as orc  # noqa: E402


def _gather(loc, shape_of, world):
    parts = [torch.empty(shape_of(r), dtype=torch.float64, device='cuda') for r in range(world)]
    dist.all_gather(parts, loc.contiguous())
    return parts


def euler_chain(nc, nsteps, transport, native=True, k=2, Tf=1e9, cfl=0.18, ic='shuosher'):
    """Returns (ok, detail) on rank 0, (None, '') elsewhere."""
    rank, world = dist.get_rank(), dist.get_world_size()
    ops = orc.make_ops(k); o = orc.operators(k)
    plug = ic_mod.EULER[ic]
    dx = (plug.xmax - plug.xmin) / nc
    x = ic_mod.gauss_points(plug.xmin, dx, nc, o['xg'])
    vals = plug.initial_condition(x.reshape(-1))
    u = np.stack([orc.project(ops, o['wg'], np.ascontiguousarray(np.asarray(v).reshape(nc, k + 1))) for v in vals])
    Mdx2 = 100 * dx * dx
    with DistributedDg1d('euler', 'rusanov', k, nc, limiter=True, char_lim=1, periodic=False, native=native, transport=transport) as dd:
        dd.dev.pack(np.ascontiguousarray(u[:, dd.start:dd.stop]).reshape(3, 1, dd.stop - dd.start, k + 1))
        if nsteps is None:
            it, t = dd.solve(cfl, dx, Tf, Mdx2)
        else:
            dd.dev.reset_control(); dd.max_speed()
            dd.run_steps(nsteps, cfl, dx, Tf, Mdx2)
            dd.check_transport()
            it, t = int(dd.dev.it[0].item()), float(dd.dev.t[0].item())
        loc = dd.dev.unpack().reshape(3, dd.stop - dd.start, k + 1)
        parts = _gather(loc, lambda r: (3,) + (block_partition(nc, world, r)[1] - block_partition(nc, world, r)[0], k + 1), world)
        used = dd.transport
    if rank != 0:
        return None, ''
    got = torch.cat(parts, 1).cpu().numpy()
    ref = u.copy()
    oit, ot, _ = orc.euler_solve(ops, 'rusanov', 1, ref, cfl, dx, Tf, Mdx2, max_steps=-1 if nsteps is None else nsteps)
    same_t = (t == ot) or (np.isnan(t) and np.isnan(ot))
    ok = bool(np.array_equal(got, ref, equal_nan=True) and it == oit and same_t)
    return ok, 'euler chain %s nc=%d world=%d transport=%s: it=%d (oracle %d) t=%r bit-identical=%s' % (ic, nc, world, used, it, oit, t, ok)


def burgers_ring(nc, nsteps, transport, native=True, k=2):
    rank, world = dist.get_rank(), dist.get_world_size()
    ops = orc.make_ops(k); o = orc.operators(k)
    dx = 1.0 / nc
    x = ic_mod.gauss_points(0.0, dx, nc, o['xg'])
    u = orc.project(ops, o['wg'], np.ascontiguousarray(ic_mod.SCALAR['sin2pi'].initial_condition(x.reshape(-1)).reshape(nc, k + 1)))
    with DistributedDg1d('burger', 'roe', k, nc, limiter=True, periodic=True, native=native, transport=transport) as dd:
        dd.dev.pack(np.ascontiguousarray(u[dd.start:dd.stop]).reshape(1, 1, dd.stop - dd.start, k + 1))
        dd.dev.reset_control(); dd.max_speed()
        dd.run_steps(nsteps, 0.18, dx, 1e9, 10 * dx * dx)
        dd.check_transport()
        loc = dd.dev.unpack().reshape(dd.stop - dd.start, k + 1)
        parts = _gather(loc, lambda r: (block_partition(nc, world, r)[1] - block_partition(nc, world, r)[0], k + 1), world)
        used = dd.transport
    if rank != 0:
        return None, ''
    got = torch.cat(parts, 0).cpu().numpy()
    ref = u.copy()
    orc.scalar_solve(ops, 'burger', 'roe', True, ref, 0.18, dx, 1e9, 10 * dx * dx, max_steps=nsteps)
    ok = bool(np.array_equal(got, ref))
    return ok, 'burgers ring nc=%d world=%d transport=%s: bit-identical=%s' % (nc, world, used, ok)


def run_all(transports=('peer', 'nccl'), native=True, log=None):
    """All cases on the given transports.  Needs an initialised NCCL process group and the CUDA device set.
    Returns {transport: {case: bool}} on rank 0 (None on the others)."""
    world = dist.get_world_size()
    out = {}
    for tr in transports:
        res = {}
        cases = [('euler_chain', lambda: euler_chain(4001, 25, tr, native)),
                 ('burgers_ring', lambda: burgers_ring(3000, 30, tr, native)),
                 # Euler tiles own 124 cells, Burgers degree 2 tiles 188: 2*TILE+1 cells per rank
                 ('euler_chain_one_cell_tail', lambda: euler_chain(249 * world, 12, tr, native, ic='sod')),
                 ('burgers_ring_one_cell_tail', lambda: burgers_ring(377 * world, 12, tr, native)),
                 ('euler_blow_up', lambda: euler_chain(200, None, tr, native, Tf=1.8, cfl=0.9 / 5))]
        for name, fn in cases:
            ok, msg = fn()
            if dist.get_rank() == 0:
                res[name] = ok
                if log is not None:
                    log(msg)
        out[tr] = res
    return out if dist.get_rank() == 0 else None


def main():
    native = '--python' not in sys.argv   # default: the C step loop (dg1d_dist_run_steps); --python: torch.distributed path
    transports = ('nccl',) if '--nccl' in sys.argv else ('peer',) if '--peer' in sys.argv else ('peer', 'nccl')
    torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
    dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
    res = run_all(transports, native, log=print)
    dist.destroy_process_group()
    if res is not None:
        print(res)
        if not all(all(v.values()) for v in res.values()):
            sys.exit(1)


if __name__ == '__main__':
    main()
