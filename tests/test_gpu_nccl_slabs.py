"""GPU suite, needs >= 2 GPUs (skipped on a single-GPU box): the row-slab decomposition over
real NCCL send/recv, one process per GPU under torchrun, bit-compared with the CPU red-black
statement; the stream-ordered and the host-synchronous NCCL exchange and the library's own peer-memory
transport (csrc/fluid_peer.cu, CUDA IPC) are exercised."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_nccl_slabs_bitwise():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs (the single-GPU emulation of the same path is tests/test_gpu_slabs.py)")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "helpers", "nccl_slab_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert out.stdout.count("bitwise equal") == 9, out.stdout
