"""512 x N=512 ensemble, a few frames, for an ncu launch list: python scripts/r02_ens_list.py <stream mode 0|1|2> <frames>"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

mode, frames = int(sys.argv[1]), int(sys.argv[2])
if os.environ.get("FLUID_ALT_LIB"):  # an experimental build of the library (scripts/r02_run18.sh)
    fl.LIB_PATH = os.path.join(ROOT, os.environ["FLUID_ALT_LIB"])
M, N = 512, 512
fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=M)
fs.SetStreamKernel(mode)
fs.SetFastDivision(2)  # proofs at once: the list shows the steady state from the first frame on
for b in range(M):
    cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
    fs.SetSources([(fl.DENS, cx, cy, fl.SRC_PREV_SET, 100.0), (fl.U, cx, cy, fl.SRC_PREV_SET, 5.0), (fl.V, cx, cy, fl.SRC_PREV_SET, 5.0)], sim=b)
fs.Update(0.1, frames)
fs.Sync()
print("ms per frame (all frames):", fs.LastStepMs() / frames)
if len(sys.argv) > 3:  # ... and the per-class times of some more
    more = int(sys.argv[3])
    fs.SetProfiling(True)
    fs.Profile(reset=True)
    fs.Update(0.1, more)
    fs.Sync()
    prof = fs.Profile(reset=True)
    print(os.environ.get("FLUID_ALT_LIB", "default"), "ms per frame: %.2f" % (fs.LastStepMs() / more), {k: round(v[0] / more, 2) for k, v in prof.items()})
fs.close()
