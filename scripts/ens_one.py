"""Ensemble of N=512 members on one GPU: ms per step and per-class times, streaming kernels on and off."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "fluid-simulation_b200"))
import fluid_b200 as fl
N = 512
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 128
for stream in (True, False):
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=1, batch=batch)
    fs.SetStreamKernel(stream)
    for b in range(batch):
        cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
        fs.SetSources([(2, cx, cy, 0, 100.0), (0, cx, cy, 0, 5.0), (1, cx, cy, 0, 5.0)], sim=b)
    fs.Update(0.1, 4); fs.Sync()
    fs.Update(0.1, 4); fs.Sync()
    fs.SetProfiling(True); fs.Profile(reset=True)
    fs.Update(0.1, 4); fs.Sync()
    prof = fs.Profile(reset=True)
    print("stream", stream, "ms/step", fs.LastStepMs() / 4, {k: round(v[0] / 4, 3) for k, v in prof.items()})
    fs.close()
