"""A few reference-order solves alone (for ncu): python scripts/r02_rowwave_one.py <N> <mode 0|2|3> <solves> [a c]"""
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

if os.environ.get("FLUID_ALT_LIB"):  # an experimental build of the library
    fl.LIB_PATH = os.path.join(ROOT, os.environ["FLUID_ALT_LIB"])

N, mode, solves = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
a = np.float32(sys.argv[4]) if len(sys.argv) > 4 else np.float32(1.0)
c = np.float32(sys.argv[5]) if len(sys.argv) > 5 else np.float32(4.0)
rng = np.random.default_rng(1)
fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=20)
fs.SetWavefrontPersistent(mode)
fs.SetFastDivision(2)
fs.upload(fl.U, rng.uniform(-1, 1, (N + 2) ** 2).astype(np.float32))
fs.upload(fl.V, rng.uniform(-1, 1, (N + 2) ** 2).astype(np.float32))
fs.lin_solve(0, fl.U, fl.V, a, c)
fs.Sync()
t0 = time.perf_counter()
for _ in range(solves):
    fs.lin_solve(0, fl.U, fl.V, a, c)
fs.Sync()
print("N=%d mode=%d a=%g c=%g: %.3f ms per solve" % (N, mode, a, c, (time.perf_counter() - t0) * 1e3 / solves))
fs.close()
