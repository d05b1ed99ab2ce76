"""Under torchrun: per-frame device time of a slab with the peer transport, one stream, class profile per 10 frames."""
import json, os, sys
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
for transport in ("peer", "nccl"):
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, device=local, slab=(rank, world, rows))
    tr = PeerExchange(fs, rank, world, local) if transport == "peer" else NcclExchange(fs, rank, world, local)
    c, cy = N // 2, rows * world // 2
    fs.SetSources([(fl.DENS, c, cy, 0, 100.0), (fl.U, c, cy, 0, 5.0), (fl.V, c, cy, 0, 5.0)])
    fs.SetOverlap(False)
    fs.SetProfiling(True)
    series, profs = [], []
    for k in range(40):
        fs.Update(0.1, 1)
        fs.Sync()
        series.append(round(fs.LastStepMs(), 3))
        if k % 10 == 9:
            p = fs.Profile(reset=True)
            profs.append({kk: round(v[0] / 10, 3) for kk, v in p.items()})
    if rank == 0:
        print(json.dumps({"transport": transport, "ms_by_frame": series, "classes_per_10_frames": profs}), flush=True)
    if transport == "peer":
        tr.close()
    fs.close()
    dist.barrier()
dist.destroy_process_group()
