"""Step time and hand-off latency of the row pipeline: one band / two bands / two sweeps, timed alone."""
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

if os.environ.get("FLUID_ALT_LIB"):  # an experimental build of the library
    fl.LIB_PATH = os.path.join(ROOT, os.environ["FLUID_ALT_LIB"])

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
rng = np.random.default_rng(1)
for a, c in ((1.0, 4.0), (10.49, 42.96)):
    for NY, iters in ((30, 1), (60, 1), (30, 2), (120, 1), (30, 4), (240, 8)):
        fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=iters, slab=(0, 1, NY))
        fs.SetWavefrontPersistent(3)
        fs.SetFastDivision(2)
        fs.upload(fl.U, rng.uniform(-1, 1, (NY + 2) * (N + 2)).astype(np.float32))
        fs.upload(fl.V, rng.uniform(-1, 1, (NY + 2) * (N + 2)).astype(np.float32))
        for _ in range(3):
            fs.lin_solve(0, fl.U, fl.V, np.float32(a), np.float32(c))
        fs.Sync()
        reps = 20
        t0 = time.perf_counter()
        for _ in range(reps):
            fs.lin_solve(0, fl.U, fl.V, np.float32(a), np.float32(c))
        fs.Sync()
        us = (time.perf_counter() - t0) * 1e6 / reps
        print("N=%d c=%g bands=%d sweeps=%d: %.1f us per solve" % (N, c, NY // 30, iters, us), flush=True)
        fs.close()
