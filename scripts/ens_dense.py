"""Ensemble lin_solve time on dense random fields (no subnormal fringe) vs the point-source state."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "fluid-simulation_b200"))
import numpy as np
import fluid_b200 as fl
N, batch = 512, 512
rng = np.random.default_rng(1)
x = rng.uniform(-1, 1, (N + 2) ** 2).astype(np.float32)
for stream in (True, False):
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=1, batch=batch)
    fs.SetStreamKernel(stream)
    fs.SetFastDivision(2)
    for b in range(batch):
        fs.upload(fl.U, x, sim=b)
        fs.upload(fl.U_PREV, x, sim=b)
    fs.Sync()
    fs.time_kernel(fl.K_LIN_SOLVE, 1)
    print("dense  stream", stream, "lin_solve ms", fs.time_kernel(fl.K_LIN_SOLVE, 2), "blur ms", fs.time_kernel(fl.K_BLUR, 2))
    fs.close()
