"""Round-2 probe 3: reference-order (wavefront) step times with the one-launch-per-solve kernel on and off,
and the per-class breakdown of the 512 x N=512 ensemble.  One JSON line per case."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402


def src_of(N):
    c = N // 2
    return [(fl.DENS, c, c, fl.SRC_PREV_SET, 100.0), (fl.U, c, c, fl.SRC_PREV_SET, 5.0), (fl.V, c, c, fl.SRC_PREV_SET, 5.0)]


what = sys.argv[1] if len(sys.argv) > 1 else "all"
if what in ("all", "wavefront"):
    for N, steps in ((1024, 10), (2048, 5), (8192, 3)):
        for persistent in (True, False):
            for tile in ((16, 32) if persistent else (0,)):
                fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_WAVEFRONT)
                fs.SetWavefrontPersistent(2 if persistent else 0)
                fs.SetWavefrontTile(tile)
                fs.SetSources(src_of(N))
                fs.Update(0.1, 2)
                fs.Sync()
                fs.Update(0.1, steps)
                fs.Sync()
                ms = fs.LastStepMs() / steps
                print(json.dumps({"case": "wavefront", "N": N, "persistent": persistent, "tile": tile, "ms_per_step": ms,
                                  "cell_updates_per_s": N * N / (ms * 1e-3)}), flush=True)
                fs.close()
if what in ("all", "ensemble"):
    M, N = 512, 512
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=M)
    for b in range(M):
        cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
        fs.SetSources([(fl.DENS, cx, cy, fl.SRC_PREV_SET, 100.0), (fl.U, cx, cy, fl.SRC_PREV_SET, 5.0), (fl.V, cx, cy, fl.SRC_PREV_SET, 5.0)], sim=b)
    fs.Update(0.1, 4)
    fs.Sync()
    for overlap in (True, False):
        fs.SetOverlap(overlap)
        fs.SetProfiling(not overlap)
        fs.Profile(reset=True)
        fs.Update(0.1, 5)
        fs.Sync()
        ms = fs.LastStepMs() / 5
        prof = fs.Profile(reset=True) if not overlap else {}
        print(json.dumps({"case": "ensemble 512x512", "overlap": overlap, "ms_per_step": ms, "cell_updates_per_s": M * N * N / (ms * 1e-3),
                          "classes_ms_per_step": {k: v[0] / 5 for k, v in prof.items()}, "launches": {k: v[1] // 5 for k, v in prof.items()}}), flush=True)
    fs.close()
if what in ("ensemble_modes",):
    # the same ensemble with the streaming kernels on automatic (redo feedback picks the sequence), forced to the scaled
    # division sequence, and off (tile / pass kernels only)
    M, N = 512, 512
    for mode in (1, 2, 0):
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=M)
        fs.SetStreamKernel(mode)
        for b in range(M):
            cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
            fs.SetSources([(fl.DENS, cx, cy, fl.SRC_PREV_SET, 100.0), (fl.U, cx, cy, fl.SRC_PREV_SET, 5.0), (fl.V, cx, cy, fl.SRC_PREV_SET, 5.0)], sim=b)
        fs.Update(0.1, 10)
        fs.Sync()
        fs.SetProfiling(True)
        fs.Profile(reset=True)
        fs.Update(0.1, 5)
        fs.Sync()
        ms = fs.LastStepMs() / 5
        prof = fs.Profile(reset=True)
        print(json.dumps({"case": "ensemble 512x512", "stream_kernel": mode, "ms_per_step": ms, "cell_updates_per_s": M * N * N / (ms * 1e-3),
                          "classes_ms_per_step": {k: v[0] / 5 for k, v in prof.items()}, "launches": {k: v[1] // 5 for k, v in prof.items()}}), flush=True)
        fs.close()
if what in ("wavefront_rows",):
    # the reference-order step: row pipeline (3) against tile tickets (2) and a launch per hyperplane (0)
    for N, steps in ((256, 20), (512, 20), (1024, 10), (1536, 6), (2048, 5), (3072, 4), (4096, 3), (8192, 2)):
        for mode in (3, 2, 0, 1):
            if mode == 2 and N > 2048:
                continue
            fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_WAVEFRONT)
            fs.SetWavefrontPersistent(mode)
            fs.SetSources(src_of(N))
            fs.Update(0.1, 3)
            fs.Sync()
            fs.Update(0.1, steps)
            fs.Sync()
            ms = fs.LastStepMs() / steps
            fs.SetProfiling(True)
            fs.Profile(reset=True)
            fs.Update(0.1, steps)
            fs.Sync()
            prof = fs.Profile(reset=True)
            print(json.dumps({"case": "wavefront", "N": N, "mode": mode, "ms_per_step": ms, "cell_updates_per_s": N * N / (ms * 1e-3),
                              "classes_ms_serial": {k: round(v[0] / steps, 3) for k, v in prof.items()}}), flush=True)
            fs.close()
