import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "fluid-simulation_b200"))
import numpy as np, fluid_b200 as fl
N=8192
fs=fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=1)
c=N//2
fs.SetSources([(2,c,c,0,100.0),(0,c,c,0,5.0),(1,c,c,0,5.0)])
fs.Update(0.1, 4); fs.Sync()
fs.SetFastDivision(2)
print(fs.time_kernel(fl.K_LIN_SOLVE, 1), fs.time_kernel(fl.K_LIN_SOLVE, 2))
