"""Advect kernels at N=8192: the gather kernel (1) against the row-window kernel (2), on the bench's own
field (source S1 after a few frames: mostly zero velocity) and on a dense smooth field whose backtraces
move up to `amp` cells.  Prints one JSON line per case."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK)
c = N // 2
fs.SetSources([(fl.DENS, c, c, fl.SRC_PREV_SET, 100.0), (fl.U, c, c, fl.SRC_PREV_SET, 5.0), (fl.V, c, c, fl.SRC_PREV_SET, 5.0)])
fs.Update(0.1, 8)
fs.Sync()


def measure(tag):
    for kernel in (1, 2):
        fs.SetAdvectKernel(kernel)
        out = {"field": tag, "N": N, "kernel": "gather" if kernel == 1 else "rows"}
        for name, which, nbytes in (("advect_velocity", fl.K_ADVECT_VEL, 16), ("advect", fl.K_ADVECT, 16)):
            fs.time_kernel(which, 2)
            ms = fs.time_kernel(which, reps)
            out[name + "_us"] = 1e3 * ms
            out[name + "_TBs"] = nbytes * N * N / (ms * 1e-3) / 1e12
        print(json.dumps(out), flush=True)


measure("S1 after 8 frames")
if "--quick" not in sys.argv:
    i = np.arange(N + 2, dtype=np.float32)
    for amp in (0.8, 3.0):
        scale = amp / (0.1 * N)
        u = (scale * np.sin(2 * np.pi * i / (N / 3.3) + 0.3)[None, :] * np.cos(2 * np.pi * i / (N / 2.1))[:, None]).astype(np.float32)
        v = (scale * np.cos(2 * np.pi * i / (N / 3.3))[None, :] * np.sin(2 * np.pi * i / (N / 2.1) + 1.1)[:, None]).astype(np.float32)
        fs.upload(fl.U, u.ravel())
        fs.upload(fl.V, v.ravel())
        fs.upload(fl.DENS, np.abs(u).ravel())
        measure("smooth dense, |dt*N*v| <= %.1f cells" % amp)
fs.close()
