"""CPU suite: the host-side halo pairing of the multi-GPU path, with gloo and
world_size 2 and 3 (torch.distributed on 127.0.0.1).  The tensors stand in for
the device blocks fluid_ctx_halo_regions() returns; the logic (who sends which
block to whom, in which order) is the code the NCCL run uses."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, rows, halo, pitch, nfields):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "fluid-simulation_b200"))
    from fluid_b200.slabs import exchange_pairs
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo_nb, hi_nb = rank > 0, rank < world - 1
        NY = rows + (halo if lo_nb else 0) + (halo if hi_nb else 0)
        j0 = rank * rows - (halo if lo_nb else 0)      # global row of local row 0
        own_lo = halo + 1 if lo_nb else 1
        blocks, fields = [], []
        for f in range(nfields):
            # every stored value = 1000*field + its GLOBAL row; halo rows start out poisoned
            a = torch.full((NY + 2, pitch), -1.0)
            for j in range(own_lo, own_lo + rows):
                a[j] = 1000.0 * f + (j + j0)
            h = halo + 1
            top = own_lo + rows
            flat = a.view(-1)
            blk = (flat[own_lo * pitch:(own_lo + h) * pitch] if lo_nb else None,
                   flat[0:h * pitch] if lo_nb else None,
                   flat[(top - h) * pitch:top * pitch] if hi_nb else None,
                   flat[top * pitch:(top + h) * pitch] if hi_nb else None)
            blocks.append(blk)
            fields.append(a)
        exchange_pairs(dist, rank, world, blocks)
        for f, a in enumerate(fields):
            for j in range(0, NY + 2):
                g = j + j0
                owned_somewhere = 1 <= g <= rows * world
                if owned_somewhere:
                    assert torch.all(a[j] == 1000.0 * f + g), (rank, f, j, a[j, 0].item())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_halo_pairing_with_gloo(world):
    mp.spawn(_worker, args=(world, _free_port(), 60, 8, 16, 2), nprocs=world, join=True)
