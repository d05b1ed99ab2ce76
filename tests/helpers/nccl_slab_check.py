"""Run under torchrun (one rank per GPU): the row-slab path over real NCCL send/recv (stream-ordered and
host-synchronous) and over the library's own peer-memory transport (pushes through CUDA IPC mappings),
checked bit for bit against the CPU red-black statement on rank 0.  Used by
tests/test_gpu_nccl_slabs.py; exits non-zero on any mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import fluid_b200 as fl  # noqa: E402
from fluid_b200.slabs import NcclExchange, PeerExchange  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fl.check(fl.lib().fluid_set_device(local))
    rows, steps = 80, 3
    N = rows * world
    rng = np.random.default_rng(5)
    init = {k: (rng.uniform(-1, 1, (N + 2) ** 2) * 0.5).astype(np.float32) for k in "uvd"}
    c = N // 2
    src = [(2, c, c, 0, 100.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0), (1, 9, rows, 0, 3.0), (0, 11, rows + 1, 0, -2.0)]
    ok = True
    for stream_ordered in (True, False, "peer"):
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, device=local, slab=(rank, world, rows))
        peer = None
        if stream_ordered == "peer":
            peer = PeerExchange(fs, rank, world, local)
        else:
            NcclExchange(fs, rank, world, local, stream_ordered=stream_ordered)
        fs.SetSources(src)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            fs.upload(fid, init[name])
        fs.Update(0.1, steps)
        fs.Sync()
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            out = np.zeros((N + 2) ** 2, np.float32)
            fs.download(fid, out=out)           # this rank's owned rows, zeros elsewhere
            t = torch.from_numpy(out).cuda()
            dist.reduce(t, 0, op=dist.ReduceOp.SUM)  # disjoint rows: the sum assembles the global field exactly
            if rank == 0:
                import oracle_py as op
                if name == "u":
                    f = {k: init[k].copy() for k in "uvd"}
                    g = {"u": f["u"], "v": f["v"], "d": f["d"]}
                    g.update({k: np.zeros((N + 2) ** 2, np.float32) for k in ("u0", "v0", "d0")})
                    op.build()
                    op.Oracle().update(N, g, src, 1e-4, 1e-4, 0.1, steps, order=op.GS_RB)
                    want = g
                got = t.cpu().numpy()
                same = got.tobytes() == want[name].tobytes()
                print("stream_ordered=%s field %s: %s" % (stream_ordered, name, "bitwise equal" if same else "MISMATCH"))
                ok &= same
        if peer is not None:
            peer.close()
        fs.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
