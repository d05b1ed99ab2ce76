"""Summary of a whole solve's task timeline (FLUID_RW_STAMPS): python scripts/r02_rowwave_stamps_summary.py <N>"""
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

N = int(sys.argv[1])
rng = np.random.default_rng(1)
fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=20)
fs.SetWavefrontPersistent(3)
fs.SetFastDivision(2)
fs.upload(fl.U, rng.uniform(-1, 1, (N + 2) ** 2).astype(np.float32))
fs.upload(fl.V, rng.uniform(-1, 1, (N + 2) ** 2).astype(np.float32))
for _ in range(2):
    fs.lin_solve(0, fl.U, fl.V, np.float32(1.0), np.float32(4.0))
fs.Sync()
fs.close()
last = open(path).read().split("solve ")[-1].splitlines()
print(last[0])
nb = int(last[0].split("nb=")[1].split()[0])
rows = np.array([[int(v) for v in ln.split()] for ln in last[1:]], dtype=np.int64)
t0 = rows[:, 1].min()
T = (rows[:, 1:] - t0) / 1e3
k, band = np.divmod(rows[:, 0], nb)
print("whole solve: %.1f us" % T[:, 3].max())
for kk in (0, 1, 10, 19):
    sel = k == kk
    order = np.argsort(band[sel])
    first = T[sel][order]
    print("sweep %2d: taken %8.1f..%8.1f  first wait over: band 0 %8.1f, band %d %8.1f (%.1f us per band); steps done: band 0 %8.1f last %8.1f; task duration median %.1f us"
          % (kk, first[:, 0].min(), first[:, 0].max(), first[0, 1], nb - 1, first[-1, 1], (first[-1, 1] - first[0, 1]) / max(nb - 1, 1),
             first[0, 2], first[-1, 2], np.median(first[:, 2] - first[:, 1])))
