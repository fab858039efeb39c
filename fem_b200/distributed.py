"""1-D domain decomposition over the GPUs of one node: one process per GPU, contiguous blocks
of cells, a two-cell halo exchanged with each neighbour before every RK stage (NCCL send/recv over
NVLink) and one all_reduce(MAX) of the CFL wave speed per time step.  Results are bit-identical to a
single-block run: the halo lets a rank recompute its neighbours' updated cell averages itself (F8).

The rank/neighbour logic and the exchange are backend-agnostic torch.distributed code, so the CPU
test-suite exercises them with gloo (world size 2); the kernels themselves need a GPU.

Batched-ensemble mode shards independent problems instead and never communicates (shard_problems).
"""
import ctypes as C
import os

import torch
import torch.distributed as dist


def block_partition(ncell, world, rank):
    """Cells [start, stop) of `rank`: contiguous blocks, the first ncell % world ranks get one more."""
    base, extra = divmod(int(ncell), int(world))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_problems(nprob, world, rank):
    """Ensemble mode: problems [start, stop) of `rank`; no communication on the data path."""
    return block_partition(nprob, world, rank)


def neighbours(world, rank, periodic):
    """(left, right) neighbour ranks or None at a physical (copy) boundary.  Periodic => ring."""
    left = rank - 1 if rank > 0 else (world - 1 if periodic else None)
    right = rank + 1 if rank < world - 1 else (0 if periodic else None)
    return left, right


def boundary_kinds(world, rank, periodic):
    """bc strings for Dg1dDevice.  A periodic mesh on several ranks is a ring of halos; on a single
    rank the kernels keep the periodic ghosts themselves."""
    if world == 1:
        return ('periodic', 'periodic') if periodic else ('copy', 'copy')
    left, right = neighbours(world, rank, periodic)
    return ('halo' if left is not None else 'copy', 'halo' if right is not None else 'copy')


class HaloExchanger:
    """Exchanges the packed edge cells with the left/right neighbour.  Buffers: send_left/right hold this
    rank's two leftmost/rightmost cells; recv_left/right receive the neighbours' facing cells."""

    def __init__(self, n, device, world, rank, periodic, group=None, dtype=torch.float64):
        self.left, self.right = neighbours(world, rank, periodic)
        self.group = group
        self.rank, self.world = rank, world
        mk = lambda: torch.zeros(n, dtype=dtype, device=device)
        self.send_left = mk() if self.left is not None else None
        self.send_right = mk() if self.right is not None else None
        self.recv_left = mk() if self.left is not None else None
        self.recv_right = mk() if self.right is not None else None

    def exchange(self):
        """Call after the send buffers are filled (same stream); returns when the receive buffers are
        valid for work queued on the current stream."""
        if self.left is None and self.right is None:
            return
        if self.left is not None and self.left == self.right:
            # two-rank ring: both neighbours are the same peer.  NCCL pairs the n-th send to a peer with that
            # peer's n-th receive (tags are ignored), gloo pairs by tag; this order satisfies both:
            # my send_left lands in the peer's recv_right, my send_right in its recv_left.
            ops = [dist.P2POp(dist.isend, self.send_left, self.left, self.group, tag=1),
                   dist.P2POp(dist.isend, self.send_right, self.right, self.group, tag=2),
                   dist.P2POp(dist.irecv, self.recv_right, self.right, self.group, tag=1),
                   dist.P2POp(dist.irecv, self.recv_left, self.left, self.group, tag=2)]
        else:
            ops = []
            if self.left is not None:
                ops.append(dist.P2POp(dist.isend, self.send_left, self.left, self.group))
                ops.append(dist.P2POp(dist.irecv, self.recv_left, self.left, self.group))
            if self.right is not None:
                ops.append(dist.P2POp(dist.isend, self.send_right, self.right, self.group))
                ops.append(dist.P2POp(dist.irecv, self.recv_right, self.right, self.group))
        for req in dist.batch_isend_irecv(ops):
            req.wait()


class DistributedDg1d:
    """Dg1dDevice for this rank's block + the per-stage halo exchange and per-step speed reduction."""

    def __init__(self, pde, flux, degree, ncell_global, limiter=False, char_lim=0, periodic=True, gamma=1.4,
                 group=None, device=None, native=True, transport='auto'):
        from .device import Dg1dDevice
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        self.periodic = periodic
        self.start, self.stop = block_partition(ncell_global, self.world, self.rank)
        self.ncg = int(ncell_global)
        self.dev = Dg1dDevice(pde, flux, degree, self.stop - self.start, limiter=limiter, char_lim=char_lim,
                              bc=boundary_kinds(self.world, self.rank, periodic), cell_offset=self.start,
                              ncell_global=ncell_global, gamma=gamma, device=device)
        self.halo = HaloExchanger(self.dev.halo_doubles(), self.dev.device, self.world, self.rank, periodic, group)
        # native=True: the step loop, the exchanges and the reduction run in C (dg1d_dist_run_steps) on a communicator
        # of the library's own; native=False keeps them in Python on torch.distributed's (same results, more overhead)
        self._dist = None
        if native and self.world > 1:
            from . import _lib
            lib = self.dev.lib
            if not lib.dg1d_dist_available():
                raise _lib.Dg1dError('libnccl.so.2 not found in the process: native multi-GPU stepping is unavailable')
            uid = C.create_string_buffer(128)
            if self.rank == 0:
                _lib.check(lib.dg1d_dist_unique_id(uid))
            box = [uid.raw]
            dist.broadcast_object_list(box, src=0, group=group)
            left, right = neighbours(self.world, self.rank, periodic)
            h = C.c_void_p()
            with torch.cuda.device(self.dev.device):
                _lib.check(lib.dg1d_dist_create(box[0], self.world, self.rank, -1 if left is None else left,
                                                -1 if right is None else right, C.byref(h)))
            self._dist = h
            n = self.dev.halo_doubles()
            self._nbuf = [torch.zeros(n, dtype=torch.float64, device=self.dev.device) for _ in range(4)]
            # transport 'peer': halo and CFL reduction as single kernels storing into the neighbours' HBM (CUDA IPC
            # mailboxes over NVLink); 'nccl': ncclSend/Recv + ncclAllReduce; 'auto': peer when every rank can map
            # every mailbox, else nccl.  The choice is collective: all ranks end up on the same transport.
            if transport not in ('auto', 'peer', 'nccl'):
                raise ValueError("transport must be 'auto', 'peer' or 'nccl'")
            self.transport = 'nccl'
            if transport != 'nccl':
                self._connect_peers(required=(transport == 'peer'))
        else:
            self.transport = 'nccl' if self.world > 1 else 'none'

    def _connect_peers(self, required):
        from . import _lib
        lib = self.dev.lib
        mine = C.create_string_buffer(64)
        with torch.cuda.device(self.dev.device):
            rc = lib.dg1d_dist_p2p_open(self._dist, self.dev.h, mine)
        fail = os.environ.get('DG1D_TEST_P2P_FAIL', '')      # test hook "open:<rank>" / "connect:<rank>": pretend that step failed there
        if fail == 'open:%d' % self.rank:
            rc = -3
        ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=self.dev.device)
        err = '' if rc == 0 else _lib.last_error()
        blob = torch.frombuffer(bytearray(mine.raw), dtype=torch.uint8).to(self.dev.device)
        allb = torch.empty(64 * self.world, dtype=torch.uint8, device=self.dev.device)
        dist.all_gather_into_tensor(allb, blob, group=self.group)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 1:
            handles = bytes(allb.cpu().numpy().tobytes())
            with torch.cuda.device(self.dev.device):
                rc = lib.dg1d_dist_p2p_connect(self._dist, handles)
            if fail == 'connect:%d' % self.rank:
                rc = -3
            err = '' if rc == 0 else _lib.last_error()
            ok.fill_(1 if rc == 0 else 0)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 1:
            self.transport = 'peer'
            dist.barrier(group=self.group)        # every mailbox is zeroed and mapped before anyone writes into one
            return
        if required:
            raise _lib.Dg1dError('peer-memory transport unavailable on some rank' + (': ' + err if err else ''))
        # stay on NCCL everywhere: a rank that did connect must not use its mapping alone
        _lib.check(lib.dg1d_dist_use_nccl(self._dist))

    def check_transport(self):
        """Synchronise the device and raise if a peer-memory wait timed out (the kernels give up after 20 s instead of
        hanging; every later exchange then bails out at once and the state is garbage).  solve() and close() call it;
        after run_steps() it is the caller's to call before trusting the state."""
        if self._dist is None or self.transport != 'peer':
            return
        from . import _lib
        n = C.c_int64(0)
        with torch.cuda.device(self.dev.device):
            _lib.check(self.dev.lib.dg1d_dist_p2p_error(self._dist, C.byref(n)))
        if n.value:
            raise _lib.Dg1dError('a peer-memory exchange timed out waiting for a neighbour rank')

    def _exchange(self, soa):
        h = self.halo
        if h.left is None and h.right is None:
            return
        self.dev.halo_pack(soa, h.send_left, h.send_right)
        h.exchange()
        self.dev.halo_unpack(soa, h.recv_left, h.recv_right)

    def max_speed(self):
        """Initial CFL reduction; also marks the ghosts stale (call it after loading a state)."""
        self.invalidate_halo()
        self.dev.max_speed()
        self._reduce_speed()

    def close(self, check=True):
        """Collective: every rank calls it.  Raises if a peer-memory exchange had timed out (the state is then garbage)."""
        if self._dist is None:
            return
        err = None
        try:
            if check:
                self.check_transport()
        except Exception as e:        # still release everything, then report
            err = e
        if self.transport == 'peer' and dist.is_initialized():
            # no rank may unmap or free a mailbox a neighbour's last kernel could still be storing into
            torch.cuda.synchronize(self.dev.device)
            dist.barrier(group=self.group)
        self.dev.lib.dg1d_dist_destroy(self._dist)
        self._dist = None
        if err is not None:
            raise err

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close(check=exc[0] is None)
        return False

    def __del__(self):
        # last resort for a forgotten close(): no collective call can be made from a finaliser, so only this rank's
        # communicator, streams and mappings are released (after its own work has drained)
        h = getattr(self, '_dist', None)
        if h is not None:
            try:
                torch.cuda.synchronize(self.dev.device)
                self.dev.lib.dg1d_dist_destroy(h)
            except Exception:
                pass
            self._dist = None

    def _reduce_speed(self):
        if self._dist is not None:
            from . import _lib
            d = self.dev
            with torch.cuda.device(d.device):
                _lib.check(d.lib.dg1d_dist_reduce_speed(self._dist, d.speed.data_ptr(), d.nprob, d._stream()))
        elif self.world > 1:
            dist.all_reduce(self.dev.speed, op=dist.ReduceOp.MAX, group=self.group)   # F9; max is exact

    def step(self, cfl, dx, Tf, Mdx2, max_steps=-1):
        """One time step driven from Python (native=False): exchange the input halo, run the stage, three times;
        then reduce the wave speed.  Same results as the C loop, which additionally overlaps the exchange with the
        interior tiles (dg1d_dist_run_steps) and is what native=True uses."""
        d = self.dev
        if self.world == 1:
            d.run_steps(1, cfl, dx, Tf, Mdx2, max_steps)
            return
        d.step_control(cfl, dx, Tf, max_steps)
        X, Y, Z = d.bufs
        self._exchange(X)
        d.stage(0, X, None, Y, Mdx2)
        self._exchange(Y)
        d.stage(1, Y, X, Z, Mdx2)
        self._exchange(Z)
        d.stage(2, Z, X, Y, Mdx2, with_speed=True)
        d.bufs = [Y, X, Z]
        self._reduce_speed()

    def invalidate_halo(self):
        """Call after changing the state from outside (pack/project): the ghosts must be exchanged again."""
        if self._dist is not None:
            self.dev.lib.dg1d_dist_invalidate_halo(self._dist)

    def run_steps(self, n, cfl, dx, Tf, Mdx2, max_steps=-1, graph=False):
        if self.world == 1:
            (self.dev.run_steps_graphed if graph else self.dev.run_steps)(n, cfl, dx, Tf, Mdx2, max_steps)
            return
        if self._dist is not None:
            from . import _lib
            d = self.dev
            ptrs = (C.c_void_p * 3)(*[b.data_ptr() for b in d.bufs])
            log_ptr, cap = (d.dt_log.data_ptr(), d.dt_log.shape[1]) if d.dt_log is not None else (None, 0)
            sl, sr, rl, rr = self._nbuf
            with torch.cuda.device(d.device):
                _lib.check(d.lib.dg1d_dist_run_steps(d.h, self._dist, int(n), ptrs, cfl, dx, Tf, int(max_steps), float(Mdx2),
                                                     d.t.data_ptr(), d.dt.data_ptr(), d.lam.data_ptr(), d.speed.data_ptr(),
                                                     d.it.data_ptr(), d.done.data_ptr(), log_ptr, cap, sl.data_ptr(),
                                                     sr.data_ptr(), rl.data_ptr(), rr.data_ptr(), d._stream()))
            by_ptr = {b.data_ptr(): b for b in d.bufs}
            d.bufs = [by_ptr[ptrs[i]] for i in range(3)]
            return
        for _ in range(n):
            self.step(cfl, dx, Tf, Mdx2, max_steps)

    def solve(self, cfl, dx, Tf, Mdx2, chunk=16):
        self.dev.reset_control()
        self.max_speed()
        while True:
            self.run_steps(chunk, cfl, dx, Tf, Mdx2)
            finished = bool(self.dev.done.all().item())      # t, dt are identical on every rank
            self.check_transport()      # a timed-out exchange leaves garbage behind: raise instead of returning it
            if finished:
                break
        return int(self.dev.it[0].item()), float(self.dev.t[0].item())
