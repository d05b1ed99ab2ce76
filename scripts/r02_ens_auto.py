"""512 x N=512 ensemble on automatic settings: ms per frame in groups of five frames (how fast the division proofs and
the redo feedback settle on the scaled sequence)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

M, N = 512, 512
fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=M)
for b in range(M):
    cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
    fs.SetSources([(fl.DENS, cx, cy, fl.SRC_PREV_SET, 100.0), (fl.U, cx, cy, fl.SRC_PREV_SET, 5.0), (fl.V, cx, cy, fl.SRC_PREV_SET, 5.0)], sim=b)
for group in range(12):
    fs.Update(0.1, 5)
    fs.Sync()
    print("frames %3d-%3d: %.2f ms per frame" % (5 * group, 5 * group + 4, fs.LastStepMs() / 5), flush=True)
fs.close()
