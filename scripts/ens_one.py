import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "fluid-simulation_b200"))
import fluid_b200 as fl
N, batch = 512, 128
fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=1, batch=batch)
fs.SetFastDivision(2)
for b in range(batch):
    cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
    fs.SetSources([(2, cx, cy, 0, 100.0), (0, cx, cy, 0, 5.0), (1, cx, cy, 0, 5.0)], sim=b)
fs.Update(0.1, 4); fs.Sync()
print(fs.LastStepMs() / 4)
