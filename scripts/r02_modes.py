"""One grid (or an ensemble) with the streaming kernels on automatic and forced to the scaled division sequence:
python scripts/r02_modes.py <N> <members> <frames>; prints ms per frame and the per-class times."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402

N, M, frames = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
for mode in (1, 2):
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=M)
    fs.SetStreamKernel(mode)
    for b in range(M):
        cx, cy = (N // 2, N // 2) if M == 1 else (1 + (37 * b) % N, 1 + (101 * b) % N)
        fs.SetSources([(fl.DENS, cx, cy, fl.SRC_PREV_SET, 100.0), (fl.U, cx, cy, fl.SRC_PREV_SET, 5.0), (fl.V, cx, cy, fl.SRC_PREV_SET, 5.0)], sim=b)
    fs.Update(0.1, 40)
    fs.Sync()
    fs.Update(0.1, frames)
    fs.Sync()
    ms = fs.LastStepMs() / frames
    fs.SetProfiling(True)
    fs.Profile(reset=True)
    fs.Update(0.1, frames)
    fs.Sync()
    prof = fs.Profile(reset=True)
    print(json.dumps({"N": N, "members": M, "stream_kernel": mode, "ms_per_frame": ms, "cell_updates_per_s": M * N * N / (ms * 1e-3),
                      "classes_ms_serial": {k: round(v[0] / frames, 3) for k, v in prof.items()}}), flush=True)
    fs.close()
