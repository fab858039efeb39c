"""N>1 host path on CPU: world_size-2 (and 3) gloo process groups.

What is covered here without a GPU: the block partition, neighbour/boundary selection, the halo exchange
pattern (chain and 2-rank ring, where both neighbours are the same peer), the per-step all_reduce(MAX), and
the protocol itself -- a two-cell halo exchanged once per RK stage, with the u^n halo kept from stage 0,
reproduces the serial result BIT FOR BIT.  The protocol test uses the CPU oracle as the
stand-in for the stage kernel on an array extended by the ghost cells."""
import os
import socket
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fem_b200.distributed import HaloExchanger, block_partition, boundary_kinds, neighbours, shard_problems


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_partition_and_neighbours():
    for nc, P in ((100, 2), (101, 4), (10 ** 8, 8), (7, 3)):
        spans = [block_partition(nc, P, r) for r in range(P)]
        assert spans[0][0] == 0 and spans[-1][1] == nc
        assert all(spans[r][1] == spans[r + 1][0] for r in range(P - 1))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1
    assert shard_problems(4096, 8, 3) == (1536, 2048)
    assert neighbours(4, 0, False) == (None, 1) and neighbours(4, 3, False) == (2, None)
    assert neighbours(4, 0, True) == (3, 1) and neighbours(2, 1, True) == (0, 0)
    assert boundary_kinds(1, 0, True) == ('periodic', 'periodic') and boundary_kinds(1, 0, False) == ('copy', 'copy')
    assert boundary_kinds(3, 0, False) == ('copy', 'halo') and boundary_kinds(3, 2, False) == ('halo', 'copy')
    assert boundary_kinds(2, 0, True) == ('halo', 'halo')


def _exchange_worker(rank, world, port, periodic, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        n = 6
        h = HaloExchanger(n, 'cpu', world, rank, periodic)
        if h.send_left is not None:
            h.send_left.copy_(torch.arange(n, dtype=torch.float64) + 100 * rank + 10)    # "my leftmost cells"
        if h.send_right is not None:
            h.send_right.copy_(torch.arange(n, dtype=torch.float64) + 100 * rank + 20)   # "my rightmost cells"
        h.exchange()
        ok = True
        if h.left is not None:    # left ghosts <- left neighbour's rightmost cells
            ok &= bool((h.recv_left == torch.arange(n, dtype=torch.float64) + 100 * h.left + 20).all())
        if h.right is not None:   # right ghosts <- right neighbour's leftmost cells
            ok &= bool((h.recv_right == torch.arange(n, dtype=torch.float64) + 100 * h.right + 10).all())
        s = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(s, op=dist.ReduceOp.MAX)
        ok &= (s.item() == world)
        open(os.path.join(out, 'ok%d' % rank), 'w').write(str(ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world,periodic', [(2, False), (2, True), (3, True), (3, False)])
def test_halo_exchange_pattern_gloo(world, periodic):
    with tempfile.TemporaryDirectory() as out:
        mp.spawn(_exchange_worker, args=(world, _free_port(), periodic, out), nprocs=world, join=True)
        for r in range(world):
            assert open(os.path.join(out, 'ok%d' % r)).read() == 'True', r


def _protocol_worker(rank, world, port, out, nsteps):
    from oracle import dg1d_oracle as orc
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        k, char_lim, flux = 2, 1, 'rusanov'
        nd = k + 1
        u_all = np.load(os.path.join(out, 'init.npy'))
        nc = u_all.shape[1]
        dx = 10.0 / nc
        cfl, Mdx2 = 0.9 / 5, 100.0 * dx * dx
        ops = orc.make_ops(k)
        a, b = block_partition(nc, world, rank)
        left, right = neighbours(world, rank, False)
        gl, gr = (2 if left is not None else 0), (2 if right is not None else 0)
        ext = np.zeros((3, gl + (b - a) + gr, nd))
        ext[:, gl:gl + b - a] = u_all[:, a:b]
        h = HaloExchanger(3 * nd * 2, 'cpu', world, rank, False)
        loc = slice(gl, gl + b - a)

        def exchange(arr):
            if h.send_left is not None:
                h.send_left.copy_(torch.from_numpy(arr[:, gl:gl + 2].reshape(-1).copy()))
            if h.send_right is not None:
                h.send_right.copy_(torch.from_numpy(arr[:, gl + b - a - 2:gl + b - a].reshape(-1).copy()))
            h.exchange()
            if h.recv_left is not None:
                arr[:, 0:2] = h.recv_left.numpy().reshape(3, 2, nd)
            if h.recv_right is not None:
                arr[:, -2:] = h.recv_right.numpy().reshape(3, 2, nd)

        speed = torch.tensor([orc.euler_max_speed(ops, np.ascontiguousarray(ext[:, loc]))], dtype=torch.float64)
        dist.all_reduce(speed, op=dist.ReduceOp.MAX)
        for _ in range(nsteps):
            dt = cfl * dx / speed.item()
            lam = dt / dx
            exchange(ext)
            u0 = ext.copy()                       # u^n including the halo received at stage 0
            for rk in range(3):
                if rk > 0:
                    exchange(ext)
                res = orc.euler_residual(ops, flux, np.ascontiguousarray(ext))
                orc.rk_update(ops, rk, u0, ext, res, lam)
                orc.euler_limiter(ops, char_lim, ext, Mdx2)
            speed = torch.tensor([orc.euler_max_speed(ops, np.ascontiguousarray(ext[:, loc]))], dtype=torch.float64)
            dist.all_reduce(speed, op=dist.ReduceOp.MAX)
        np.save(os.path.join(out, 'final%d.npy' % rank), ext[:, loc])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_two_cell_halo_protocol_is_bitwise_serial(world):
    """Shu-Osher, degree 2, characteristic TVB limiter (config 4 at small size), split over `world` ranks."""
    from oracle import dg1d_oracle as orc
    from fem_b200 import ic as ic_mod
    nc, k, nsteps = 240, 2, 6
    ops = orc.make_ops(k)
    o = orc.operators(k)
    dx = 10.0 / nc
    x = ic_mod.gauss_points(-5.0, dx, nc, o['xg'])
    vals = ic_mod.EULER['shuosher'].initial_condition(x.reshape(-1))
    u = np.stack([orc.project(ops, o['wg'], np.ascontiguousarray(np.asarray(v).reshape(nc, k + 1))) for v in vals])
    # advance a little first so that the limiter is active near the shock and the split points see waves
    orc.euler_solve(ops, 'rusanov', 1, u, 0.18, dx, 1e9, 100 * dx * dx, max_steps=30)
    with tempfile.TemporaryDirectory() as out:
        np.save(os.path.join(out, 'init.npy'), u)
        mp.spawn(_protocol_worker, args=(world, _free_port(), out, nsteps), nprocs=world, join=True)
        got = np.concatenate([np.load(os.path.join(out, 'final%d.npy' % r)) for r in range(world)], axis=1)
    serial = u.copy()
    orc.euler_solve(ops, 'rusanov', 1, serial, 0.18, dx, 1e9, 100 * dx * dx, max_steps=nsteps)
    assert np.array_equal(got, serial)
