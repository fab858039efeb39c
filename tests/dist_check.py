#!/usr/bin/env python3
"""Multi-GPU parity check, run by hand under torchrun (not collected by pytest: it needs N GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dist_check.py

Every rank owns a block of the mesh (NCCL halo exchange per stage, all_reduce(MAX) per step); the gathered
result must be bit-identical to the CPU oracle's serial run."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import ic as ic_mod  # noqa: E402
from fem_b200.distributed import DistributedDg1d  # noqa: E402
from oracle import dg1d_oracle as orc  # noqa: E402


NATIVE = '--python' not in sys.argv   # default: the C step loop (dg1d_dist_run_steps); --python: torch.distributed path
TRANSPORT = 'nccl' if '--nccl' in sys.argv else 'peer' if '--peer' in sys.argv else 'auto'


def main():
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
    dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
    ok = True
    # ---- Euler Shu-Osher k=2 characteristic limiter (config 4 at small size), chain with copy BC at the ends
    nc, k, nsteps = 4001, 2, 25
    ops = orc.make_ops(k); o = orc.operators(k)
    dx = 10.0 / nc
    x = ic_mod.gauss_points(-5.0, dx, nc, o['xg'])
    vals = ic_mod.EULER['shuosher'].initial_condition(x.reshape(-1))
    u = np.stack([orc.project(ops, o['wg'], np.ascontiguousarray(np.asarray(v).reshape(nc, k + 1))) for v in vals])
    dd = DistributedDg1d('euler', 'rusanov', k, nc, limiter=True, char_lim=1, periodic=False, native=NATIVE, transport=TRANSPORT)
    dd.dev.pack(np.ascontiguousarray(u[:, dd.start:dd.stop]).reshape(3, 1, dd.stop - dd.start, k + 1))
    dd.dev.reset_control(); dd.max_speed()
    dd.run_steps(nsteps, 0.18, dx, 1e9, 100 * dx * dx)
    loc = dd.dev.unpack().reshape(3, dd.stop - dd.start, k + 1)
    parts = [torch.empty((3, b - a, k + 1), dtype=torch.float64, device='cuda')
             for a, b in (__import__('fem_b200.distributed', fromlist=['x']).block_partition(nc, world, r) for r in range(world))]
    dist.all_gather(parts, loc.contiguous())
    if rank == 0:
        got = torch.cat(parts, 1).cpu().numpy()
        ref = u.copy()
        orc.euler_solve(ops, 'rusanov', 1, ref, 0.18, dx, 1e9, 100 * dx * dx, max_steps=nsteps)
        same = np.array_equal(got, ref)
        print('euler chain  world=%d transport=%s: bit-identical to serial oracle: %s' % (world, dd.transport, same)); ok &= same
    dd.check_transport()
    dd.close()
    # ---- Burgers k=2 TVB, periodic ring
    nc, k, nsteps = 3000, 2, 30
    ops = orc.make_ops(k); o = orc.operators(k)
    dx = 1.0 / nc
    x = ic_mod.gauss_points(0.0, dx, nc, o['xg'])
    u = orc.project(ops, o['wg'], np.ascontiguousarray(ic_mod.SCALAR['sin2pi'].initial_condition(x.reshape(-1)).reshape(nc, k + 1)))
    dd = DistributedDg1d('burger', 'roe', k, nc, limiter=True, periodic=True, native=NATIVE, transport=TRANSPORT)
    dd.dev.pack(np.ascontiguousarray(u[dd.start:dd.stop]).reshape(1, 1, dd.stop - dd.start, k + 1))
    dd.dev.reset_control(); dd.max_speed()
    dd.run_steps(nsteps, 0.18, dx, 1e9, 10 * dx * dx)
    loc = dd.dev.unpack().reshape(dd.stop - dd.start, k + 1)
    from fem_b200.distributed import block_partition
    parts = [torch.empty((b - a, k + 1), dtype=torch.float64, device='cuda') for a, b in (block_partition(nc, world, r) for r in range(world))]
    dist.all_gather(parts, loc.contiguous())
    if rank == 0:
        got = torch.cat(parts, 0).cpu().numpy()
        ref = u.copy()
        orc.scalar_solve(ops, 'burger', 'roe', True, ref, 0.18, dx, 1e9, 10 * dx * dx, max_steps=nsteps)
        same = np.array_equal(got, ref)
        print('burgers ring world=%d transport=%s: bit-identical to serial oracle: %s' % (world, dd.transport, same)); ok &= same
    dd.check_transport()
    dd.close()
    dist.destroy_process_group()
    if rank == 0 and not ok:
        sys.exit(1)


if __name__ == '__main__':
    main()
