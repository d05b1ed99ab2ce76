import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "fluid-simulation_b200"))
import fluid_b200 as fl
for N, steps in ((256, 20), (1024, 10), (2048, 5), (8192, 2)):
    for T in (16, 32, 64):
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=0)
        fs.SetWavefrontTile(T)
        c = N // 2
        fs.SetSources([(2, c, c, 0, 100.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0)])
        fs.Update(0.1, 1); fs.Sync()
        fs.Update(0.1, steps); fs.Sync()
        print("N=%d T=%d: %.3f ms/step" % (N, T, fs.LastStepMs() / steps))
        fs.close()
