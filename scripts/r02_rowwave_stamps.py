"""Timeline of the row pipeline's tasks from FLUID_RW_STAMPS: python scripts/r02_rowwave_stamps.py <N> <NY> <sweeps>"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
path = "/tmp/rw_stamps.txt"
if os.path.exists(path):
    os.remove(path)
os.environ["FLUID_RW_STAMPS"] = path
import fluid_b200 as fl  # noqa: E402

N, NY, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(1)
fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=iters, slab=(0, 1, NY))
fs.SetWavefrontPersistent(3)
fs.SetFastDivision(2)
fs.upload(fl.U, rng.uniform(-1, 1, (NY + 2) * (N + 2)).astype(np.float32))
fs.upload(fl.V, rng.uniform(-1, 1, (NY + 2) * (N + 2)).astype(np.float32))
for _ in range(3):
    fs.lin_solve(0, fl.U, fl.V, np.float32(1.0), np.float32(4.0))
fs.Sync()
fs.close()
solves = open(path).read().split("solve ")[1:]
last = solves[-1].splitlines()
print(last[0])
rows = np.array([[int(v) for v in ln.split()] for ln in last[1:]], dtype=np.int64)
t0 = rows[:, 1].min()
nb = int(last[0].split("nb=")[1].split()[0])
print("task (sweep, band): taken, first wait over, steps done, published  [us from the first task]")
for r in rows:
    k, band = divmod(int(r[0]), nb)
    print("  (%2d, %3d): %9.1f %9.1f %9.1f %9.1f" % (k, band, *[(v - t0) / 1e3 for v in r[1:]]))
