"""One band, one sweep of the row pipeline (for ncu): python scripts/r02_rowwave_band.py <N> <NY> <sweeps> [a c]"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

N, NY, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
a = np.float32(sys.argv[4]) if len(sys.argv) > 4 else np.float32(1.0)
c = np.float32(sys.argv[5]) if len(sys.argv) > 5 else np.float32(4.0)
rng = np.random.default_rng(1)
B = int(os.environ.get("RW_MEMBERS", "1"))
fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=iters, slab=(0, 1, NY), batch=B)
fs.SetWavefrontPersistent(3)
fs.SetFastDivision(2)
fs.upload(fl.U, rng.uniform(-1, 1, (NY + 2) * (N + 2)).astype(np.float32))
fs.upload(fl.V, rng.uniform(-1, 1, (NY + 2) * (N + 2)).astype(np.float32))
for _ in range(3):
    fs.lin_solve(0, fl.U, fl.V, a, c)
fs.Sync()
fs.close()
