"""Diagnostic (run under torchrun, 2+ ranks): ms/step of a row slab with each transport, overlap on and off,
and with a do-nothing exchange (pure slab compute; results wrong, timing only)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402
from fluid_b200.slabs import NcclExchange, PeerExchange  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
fl.check(fl.lib().fluid_set_device(local))
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
rows = int(sys.argv[2]) if len(sys.argv) > 2 else N
steps = 10
for transport in ("none", "peer", "nccl"):
    for overlap in (True, False):
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, device=local, slab=(rank, world, rows))
        tr = None
        if transport == "peer":
            tr = PeerExchange(fs, rank, world, local)
        elif transport == "nccl":
            tr = NcclExchange(fs, rank, world, local)
        else:
            fs.SetExchange(lambda fields: None)
            fs.SetStream(None, True)
        c, cy = N // 2, rows * world // 2
        fs.SetSources([(fl.DENS, c, cy, 0, 100.0), (fl.U, c, cy, 0, 5.0), (fl.V, c, cy, 0, 5.0)])
        fs.SetOverlap(overlap)
        fs.Update(0.1, 4)
        fs.Sync()
        dist.barrier()
        torch.cuda.synchronize()
        fs.Update(0.1, steps)
        fs.Sync()
        ms = fs.LastStepMs() / steps
        t = torch.tensor([ms], device="cuda")
        allms = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allms, t)
        if rank == 0:
            print(json.dumps({"transport": transport, "overlap": overlap, "N": N, "rows": rows, "ms_per_step_by_rank": [float(x.item()) for x in allms]}), flush=True)
        if transport == "peer":
            tr.close()
        fs.close()
        dist.barrier()
dist.destroy_process_group()
