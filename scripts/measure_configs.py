"""Time every BASELINE.json config on one GPU through the resident context (CUDA events on the
context's stream) and print one JSON object.  Not the headline bench (bench.py): a survey of the
other configurations, recorded under profiles/."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
import fluid_b200 as fl  # noqa: E402


def run(N, steps, visc, diff, order, batch=1, warm=3, varied_sources=False):
    fs = fl.FluidSystem().Setup(N, visc, diff, gs_order=order, batch=batch)
    c = N // 2
    if varied_sources:
        for b in range(batch):
            cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
            fs.SetSources([(2, cx, cy, 0, 100.0), (0, cx, cy, 0, 5.0), (1, cx, cy, 0, 5.0)], sim=b)
    else:
        fs.SetSources([(2, c, c, 0, 100.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0)])
    fs.Update(0.1, warm)
    fs.Sync()
    t0 = time.perf_counter()
    fs.Update(0.1, steps)
    fs.Sync()
    wall = time.perf_counter() - t0
    ms = fs.LastStepMs()
    fs.close()
    return {"N": N, "batch": batch, "steps": steps, "visc": visc, "diff": diff, "order": "redblack" if order else "wavefront",
            "ms_per_step": ms / steps, "wall_ms_per_step": 1e3 * wall / steps,
            "cell_updates_per_s": float(N) * N * batch * steps / (ms * 1e-3)}


out = {
    "config1_n64_diff0_1000steps_wavefront": run(64, 1000, 0.0, 0.0, 0),
    "config1_n64_diff0_1000steps_redblack": run(64, 1000, 0.0, 0.0, 1),
    "config2_n1024_100steps_wavefront": run(1024, 100, 1e-4, 1e-4, 0),
    "config2_n1024_100steps_redblack": run(1024, 100, 1e-4, 1e-4, 1),
    "config3_n8192_redblack": run(8192, 10, 1e-4, 1e-4, 1),
    "config3_n8192_wavefront": run(8192, 2, 1e-4, 1e-4, 0, warm=1),
    "config4_n16384_one_slab_redblack": run(16384, 4, 1e-4, 1e-4, 1, warm=2),
    "config5_ensemble_512x512_redblack": run(512, 10, 1e-4, 1e-4, 1, batch=512, varied_sources=True),
    "config5_ensemble_64x512_redblack": run(512, 10, 1e-4, 1e-4, 1, batch=64, varied_sources=True),
}
print(json.dumps(out, indent=1))
