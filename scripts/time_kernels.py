import sys; sys.path.insert(0,"fluid-simulation_b200")
import numpy as np, fluid_b200 as fl
N=8192
fs=fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=1)
c=N//2
fs.SetSources([(2,c,c,0,100.0),(0,c,c,0,5.0),(1,c,c,0,5.0)])
fs.Update(0.1, 6); fs.Sync()
for mode in (0,2,0,2):
    fs.SetFastDivision(mode)
    fs.time_kernel(fl.K_LIN_SOLVE, 1)
    print("S1 field  mode", mode, "lin_solve ms", fs.time_kernel(fl.K_LIN_SOLVE, 3), "blur", fs.time_kernel(fl.K_BLUR, 2))
rng=np.random.default_rng(0)
for f in (fl.U, fl.V, fl.DENS):
    fs.upload(f, rng.uniform(-1,1,(N+2)**2).astype(np.float32))
for mode in (0,2,0,2):
    fs.SetFastDivision(mode)
    fs.time_kernel(fl.K_LIN_SOLVE, 1)
    print("random field mode", mode, "lin_solve ms", fs.time_kernel(fl.K_LIN_SOLVE, 3), "blur", fs.time_kernel(fl.K_BLUR, 2))
