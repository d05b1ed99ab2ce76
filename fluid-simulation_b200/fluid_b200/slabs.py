"""Host-side transports for row slabs (multi-GPU).

The C library decides WHEN halos are exchanged (fluid_ctx_set_exchange); this
module supplies HOW, twice:

* ``NcclExchange``  -- one process per GPU (torchrun): torch.distributed
  send/recv over NCCL/NVLink between neighbouring ranks.
* ``LocalSlabGroup`` -- several slabs inside ONE process on one device, moved
  with device copies.  It emulates the ranks (one host thread per slab, a
  barrier at every exchange point) so that the decomposition can be tested on a
  single GPU.

``exchange_pairs`` is the pairing logic both share; it works on any torch
tensors, so the CPU suite runs it with gloo and world_size 2.

torch is plumbing here (process group, NCCL); no field arithmetic happens in Python.
"""
import threading

from . import FluidSystem


class _DeviceBlock:
    """A device address range as something torch.as_tensor can wrap without copying."""

    def __init__(self, ptr, nfloats):
        self.__cuda_array_interface__ = {"shape": (nfloats,), "typestr": "<f4", "data": (ptr, False), "version": 3}


def device_tensor(ptr, nfloats, device):
    import torch
    return torch.as_tensor(_DeviceBlock(ptr, nfloats), device=device)


def exchange_pairs(dist, rank, world, blocks):
    """blocks: list of (send_down, recv_down, send_up, recv_up) tensors (None where there is
    no neighbour), one tuple per field, in the same order on every rank.  Rank r's
    send_up fills rank r+1's recv_down and vice versa.  Blocking."""
    ops = []
    for send_down, recv_down, send_up, recv_up in blocks:
        if rank + 1 < world:
            ops.append(dist.P2POp(dist.isend, send_up, rank + 1))
            ops.append(dist.P2POp(dist.irecv, recv_up, rank + 1))
        if rank > 0:
            ops.append(dist.P2POp(dist.isend, send_down, rank - 1))
            ops.append(dist.P2POp(dist.irecv, recv_down, rank - 1))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


class NcclExchange:
    """Halo transport for one slab per process: torch.distributed (NCCL) send/recv.

    stream_ordered=True (default): the context runs on torch's current stream and the
    sends/receives are enqueued in stream order (torch makes the NCCL stream wait for the
    current stream, and ``wait()`` makes the current stream wait for NCCL), so a step is
    one uninterrupted stream of device work.  False: the library drains its stream before
    every exchange and the transport blocks the host until the halos have arrived."""

    def __init__(self, fs, rank, world, device, stream_ordered=True):
        import torch
        import torch.distributed as dist
        self.fs, self.rank, self.world, self.dist, self.torch = fs, rank, world, dist, torch
        self.device = torch.device("cuda", device)
        self.stream_ordered = stream_ordered
        self.cache = {}
        self.calls = 0
        self.bytes = 0
        if stream_ordered:
            self.stream = torch.cuda.Stream(device=self.device)
            fs.SetStream(self.stream.cuda_stream, True)
        fs.SetExchange(self)

    def _blocks(self, field):
        # asked afresh every time: a red-black solve may move a field to another device array
        sd, rd, su, ru, n = self.fs.HaloRegions(field)
        key = (sd, rd, su, ru, n)
        if key not in self.cache:
            self.cache[key] = tuple(device_tensor(p, n, self.device) if p else None for p in (sd, rd, su, ru))
        return self.cache[key]

    def __call__(self, fields):
        blocks = [self._blocks(f) for f in fields]
        if self.stream_ordered:
            with self.torch.cuda.stream(self.stream):
                exchange_pairs(self.dist, self.rank, self.world, blocks)  # wait() = stream-level wait on CUDA tensors
        else:
            exchange_pairs(self.dist, self.rank, self.world, blocks)
            self.torch.cuda.synchronize(self.device)
        self.calls += 1
        self.bytes += sum(b.numel() * 4 for blk in blocks for b in (blk[0], blk[2]) if b is not None)


class PeerExchange:
    """Halo transport for one slab per process WITHOUT a per-refresh callback: the library pushes halo rows straight
    into the neighbours' memory over NVLink (csrc/fluid_peer.cu).  This class only does the one-time set-up: every
    rank exports its inbox block as a CUDA IPC handle, the handles travel over torch.distributed, neighbours map
    each other's blocks.  torch is plumbing for the handshake only."""

    def __init__(self, fs, rank, world, device):
        import torch.distributed as dist
        from . import ipc_close, ipc_export, ipc_import
        self.fs, self.rank, self.world, self.dist = fs, rank, world, dist
        self.calls = self.bytes = 0  # (refreshes are counted by the library's profile, not here)
        block, nbytes = fs.PeerHaloLocal()
        import torch
        handles = [None] * world
        self.lower = self.upper = None
        err = None
        try:
            mine = ipc_export(block)
        except Exception as e:  # e.g. a platform without CUDA IPC
            mine, err = None, e
        dist.all_gather_object(handles, mine)
        try:
            if err is None and all(h is not None for h in handles):
                self.lower = ipc_import(handles[rank - 1]) if rank > 0 else None
                self.upper = ipc_import(handles[rank + 1]) if rank + 1 < world else None
            elif err is None:
                err = RuntimeError("a neighbour could not export its block")
        except Exception as e:
            err = e
        # all ranks agree on success before anybody connects (a slab that waits for a neighbour that gave up would hang)
        ok = torch.tensor([0 if err else 1], device=torch.device("cuda", device))
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            for p in (self.lower, self.upper):
                if p:
                    ipc_close(p)
            self.lower = self.upper = None
            raise RuntimeError("peer-memory halo transport unavailable on some rank (%s)" % (err or "see the other ranks"))
        fs.PeerHaloConnect(self.lower, self.upper)
        dist.barrier()  # everybody connected before anybody steps

    def close(self):
        """Call on every rank BEFORE the contexts are destroyed: nobody may still be pushing into a block that goes away."""
        from . import ipc_close
        self.fs.Sync()
        self.dist.barrier()
        for p in (self.lower, self.upper):
            if p:
                ipc_close(p)
        self.lower = self.upper = None


class LocalSlabGroup:
    """`count` slabs of `rows` rows each in one process on one device: the multi-GPU
    code path on a single GPU.  Every slab steps in its own host thread (the
    library's step call releases the GIL); exchange points are rendezvous."""

    def __init__(self, N, count, rows, device=0, slab_halo=0, **setup):
        import torch
        self.torch = torch
        self.N, self.count, self.rows = N, count, rows
        self.device = torch.device("cuda", device)
        self.slabs = [FluidSystem().Setup(N, device=device, slab=(k, count, rows), slab_halo=slab_halo, **setup)
                      for k in range(count)]
        self.barrier = threading.Barrier(count)
        self.exchanges = 0
        for k, fs in enumerate(self.slabs):
            fs.SetExchange(lambda fields, k=k: self._exchange(k, fields))

    def _exchange(self, k, fields):
        self.barrier.wait()  # every slab's stream is idle and its owned rows are final
        if k == 0:
            for f in fields:
                reg = [fs.HaloRegions(f) for fs in self.slabs]
                for r in range(self.count - 1):
                    n = reg[r][4]
                    lower_send_up, lower_recv_up = reg[r][2], reg[r][3]
                    upper_send_down, upper_recv_down = reg[r + 1][0], reg[r + 1][1]
                    device_tensor(upper_recv_down, n, self.device).copy_(device_tensor(lower_send_up, n, self.device))
                    device_tensor(lower_recv_up, n, self.device).copy_(device_tensor(upper_send_down, n, self.device))
            self.torch.cuda.synchronize(self.device)
            self.exchanges += 1
        self.barrier.wait()

    def SetSources(self, sources):
        for fs in self.slabs:
            fs.SetSources(sources)  # global coordinates; each slab keeps the entries it stores

    def upload(self, field, host_global):
        for fs in self.slabs:
            fs.upload(field, host_global)

    def download(self, field, out):
        for fs in self.slabs:
            fs.download(field, out=out)
        return out

    def Update(self, dt, nsteps=1):
        errors = []

        def run(fs):
            try:
                fs.Update(dt, nsteps)
                fs.Sync()
            except Exception as e:  # a failed slab must not leave the others waiting forever
                errors.append(e)
                self.barrier.abort()

        threads = [threading.Thread(target=run, args=(fs,)) for fs in self.slabs]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            self.barrier.reset()
            raise errors[0]

    def close(self):
        for fs in self.slabs:
            fs.close()
