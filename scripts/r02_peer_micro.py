"""Under torchrun: time halo refreshes alone (no compute between them) for both transports."""
import json
import os
import sys
import time

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
N = rows = 8192
for transport in ("peer", "nccl", "peer"):
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, device=local, slab=(rank, world, rows))
    stream = torch.cuda.Stream(device=local)
    if transport == "peer":
        fs.SetStream(stream.cuda_stream, True)
        tr = PeerExchange(fs, rank, world, local)
    else:
        tr = NcclExchange(fs, rank, world, local)
        stream = tr.stream
    for reps, gap_ms in ((20, 0.0), (200, 0.0), (50, 1.0), (50, 0.0)):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        per = []
        with torch.cuda.stream(stream):
            e0.record(stream)
            for k in range(reps):
                if gap_ms:
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    torch.cuda._sleep(int(gap_ms * 1.9e6))
                    a.record(stream)
                fs.HaloRefresh([fl.U, fl.V])
                if gap_ms:
                    b.record(stream)
                    per.append((a, b))
            e1.record(stream)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        us = 1e3 * e0.elapsed_time(e1) / reps
        inner = sum(a.elapsed_time(b) for a, b in per) * 1e3 / max(1, len(per)) if per else None
        if rank == 0:
            print(json.dumps({"transport": transport, "reps": reps, "gap_ms": gap_ms, "us_per_refresh_incl_gap": us, "us_per_refresh_alone": inner,
                              "host_wall_us_per_refresh": 1e6 * wall / reps, "blocks_env": os.environ.get("FLUID_PEER_BLOCKS")}), flush=True)
    if transport == "peer":
        tr.close()
    fs.close()
    dist.barrier()
dist.destroy_process_group()
