"""Generate tests/golden/* by running the UNMODIFIED reference solver
(oracle/_ref/libfluid_ref.so, compiled from /root/reference by oracle/Makefile).

Run in the build container only (needs /root/reference):
    make -C oracle && python tests/golden/make_golden.py

Outputs (committed):
  golden_hashes.json   FNV-1a-64 of dens/u/v after whole-frame runs (FluidSystem::Update)
  golden_fields.npz    small (N=12,16) input/output arrays for every primitive
The reference ships no golden vectors of its own, so these
ARE the pins: oracle and CUDA results are compared with them bit for bit.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import oracle_py as op  # noqa: E402

R = op.Ref()
rng = np.random.default_rng(20261106)


def rnd(N, scale=1.0):
    return (rng.uniform(-1, 1, (N + 2) ** 2) * scale).astype(np.float32)


def frames():
    out = []
    cases = [
        # N, steps, visc, diff, vorticity, damp, source
        (64, 100, 1e-4, 1e-4, 0.0, 0.0, "stam"),
        (64, 20, 1e-4, 1e-4, 0.55, 0.1, "stam"),
        (64, 1000, 0.0, 0.0, 0.0, 0.0, "stam"),      # BASELINE config 1
        (64, 30, 1e-4, 1e-4, 0.0, 0.0, "blower"),
        (100, 10, 1e-4, 1e-6, 0.0, 0.0, "stam"),
        (37, 25, 1e-3, 1e-4, 0.0, 0.0, "stam"),
        (256, 5, 1e-4, 1e-4, 0.0, 0.0, "stam"),
        (1024, 1, 1e-4, 1e-4, 0.0, 0.0, "stam"),     # BASELINE config 2, first step
        (1024, 100, 1e-4, 1e-4, 0.0, 0.0, "stam"),   # BASELINE config 2 as written: after 100 steps (~4.5 min of one core)
    ]
    for N, steps, visc, diff, vort, damp, srcname in cases:
        sim = R.sim(N, visc, diff, vort, 20, damp)
        src = op.stam_sources(N) if srcname == "stam" else op.blower_sources(N)
        sim.set_sources(src)
        sim.update(0.1, steps)
        out.append(dict(N=N, steps=steps, visc=visc, diff=diff, vorticity=vort, damp=damp, source=srcname, dt=0.1,
                        dens=op.fnv1a64(sim.field(2)), u=op.fnv1a64(sim.field(0)), v=op.fnv1a64(sim.field(1)),
                        mass=sim.total_inner_mass()))
        sim.close()
        print(out[-1])
    return out


def bordered_sources(N):
    c = N // 2
    return op.stam_sources(N) + [(0, c + 3, c - 2, 0, 4.0), (1, 3, 3, 0, -6.0), (2, c - N // 4, c, 1, 0.8), (0, 5, N - 3, 2, 0.7)]


def bordered_frames():
    """FSType::ORIGINAL_BORDERED (VelStep1B + DensStep1BS), with the shipped scene's SquareBorder and its fluid
    coefficients (SystemManager.cpp:103-108: diff 1e-6, visc 1e-4; N=192, vorticity 0.55, damp 0.1)."""
    out = []
    for N, steps, square, vort, damp in ((64, 100, True, 0.0, 0.0), (192, 10, True, 0.55, 0.1), (100, 25, False, 0.0, 0.0)):
        sim = R.sim(N, 1e-4, 1e-6, vort, 20, damp, fstype=1, square_border=square)
        sim.set_sources(bordered_sources(N))
        sim.update(0.1, steps)
        out.append(dict(N=N, steps=steps, visc=1e-4, diff=1e-6, vorticity=vort, damp=damp, square_border=square, dt=0.1, fstype=1,
                        dens=op.fnv1a64(sim.field(2)), u=op.fnv1a64(sim.field(0)), v=op.fnv1a64(sim.field(1))))
        sim.close()
        print(out[-1])
    return out


def mf_frames():
    """FSType::ORIGINAL_BORDERED_MF (VelStep1BSO + DensStep1BS) with the shipped scene's two fluids (SystemManager.cpp:99-108)."""
    fluids = [(0.000001, 0.0001), (0.0001, 0.0003)]
    out = []
    for N, steps, square, vort, damp in ((192, 10, True, 0.55, 0.1), (64, 60, True, 0.0, 0.0)):
        c = N // 2
        src = [(2, c, c, 0, 100.0), (3, c - 5, c + 4, 0, 60.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0), (3, 4, 4, 1, 0.9), (1, 3, N - 2, 0, -4.0)]
        sim = R.sim(N, 0, 0, vort, 20, damp, fluids=fluids, fstype=2, square_border=square)
        sim.set_sources(src)
        sim.update(0.1, steps)
        out.append(dict(N=N, steps=steps, vorticity=vort, damp=damp, square_border=square, dt=0.1, fstype=2, fluids=fluids,
                        dens0=op.fnv1a64(sim.density(0)), dens1=op.fnv1a64(sim.density(1)), u=op.fnv1a64(sim.field(0)), v=op.fnv1a64(sim.field(1))))
        sim.close()
    return out


def fields():
    z = {}
    for N in (12, 16):
        t = "N%d_" % N
        x, s = rnd(N), rnd(N)
        z[t + "add_source_in_x"], z[t + "add_source_in_s"] = x.copy(), s
        R.L.ref_add_source(N, x, s, 0.1)
        z[t + "add_source_out"] = x
        for b in (0, 1, 2, 3):
            x = rnd(N)
            z[t + "set_bnd_b%d_in" % b] = x.copy()
            R.L.ref_set_bnd(N, b, x)
            z[t + "set_bnd_b%d_out" % b] = x
        for b in (0, 1, 2):
            for tag, a in (("a0", 0.0), ("a1", 1.0), ("ad", 0.1 * 1e-4 * N * N)):
                c = 4.0 if tag == "a1" else 1 + 4 * a
                x, x0 = rnd(N), rnd(N)
                z[t + "lin_solve_b%d_%s_in_x" % (b, tag)], z[t + "lin_solve_b%d_%s_in_x0" % (b, tag)] = x.copy(), x0
                R.L.ref_lin_solve(N, b, x, x0, a, c, 20)
                z[t + "lin_solve_b%d_%s_out" % (b, tag)] = x
            for tag, diff in (("d0", 0.0), ("d1", 1e-3)):
                x, x0 = rnd(N), rnd(N)
                z[t + "blur_b%d_%s_in_x" % (b, tag)], z[t + "blur_b%d_%s_in_x0" % (b, tag)] = x.copy(), x0.copy()
                R.L.ref_diffuse_blur(N, b, x, x0, diff, 0.1, 20)
                z[t + "blur_b%d_%s_out_x" % (b, tag)], z[t + "blur_b%d_%s_out_x0" % (b, tag)] = x, x0
            for tag, sc in (("slow", 0.05), ("fast", 20.0)):
                d, d0, u, v = rnd(N), rnd(N), rnd(N, sc), rnd(N, sc)
                for k, a_ in (("d", d), ("d0", d0), ("u", u), ("v", v)):
                    z[t + "advect_b%d_%s_in_%s" % (b, tag, k)] = a_.copy()
                dm, de = d.copy(), d.copy()
                R.L.ref_advect(N, b, dm, d0, u, v, 0.1)
                R.L.ref_advect_euler(N, b, de, d0, u, v, 0.1)
                z[t + "advect_b%d_%s_out_midpoint" % (b, tag)], z[t + "advect_b%d_%s_out_euler" % (b, tag)] = dm, de
        u, v, p, dv = rnd(N), rnd(N), rnd(N), rnd(N)
        for k, a_ in (("u", u), ("v", v), ("p", p), ("div", dv)):
            z[t + "project_in_" + k] = a_.copy()
        R.L.ref_project(N, u, v, p, dv, 20)
        for k, a_ in (("u", u), ("v", v), ("p", p), ("div", dv)):
            z[t + "project_out_" + k] = a_
        # vel_step / dens_step through a live FluidSystem (private members via the C API)
        sim = R.sim(N, 1e-3, 1e-3, 0.0, 20, 0.0)
        ins = [rnd(N) for _ in range(6)]
        for w in range(6):
            sim.field(w)[:] = ins[w]
            z[t + "step_in_%d" % w] = ins[w]
        sim.vel_step(0.1)
        sim.dens_step(0.1)
        for w in range(6):
            z[t + "step_out_%d" % w] = sim.field(w).copy()
        sim.close()
    return z


def multifluid():
    """Two fluids with the shipped scene's properties (SystemManager.cpp:103-106) on the ORIGINAL
    pipeline: visc_avg couples them through the velocity step, each diffuses with its own coefficient."""
    N, steps = 48, 12
    fluids = [(0.000001, 0.0001), (0.0001, 0.0003)]  # (diff, visc)
    c = N // 2
    # source field ids: 0=u 1=v 2=fluid 0 3=fluid 1
    src = [(2, c, c, 0, 100.0), (3, c - 7, c + 5, 0, 60.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0), (3, 4, 4, 1, 0.9)]
    sim = R.sim(N, 0, 0, 0.0, 20, 0.0, fluids=fluids)
    sim.set_sources(src)
    sim.update(0.1, steps)
    z = dict(N=np.int32(N), steps=np.int32(steps), fluids=np.array(fluids, np.float32), sources=np.array(src, np.float64),
             u=sim.field(0).copy(), v=sim.field(1).copy(), d0=sim.density(0).copy(), d1=sim.density(1).copy())
    sim.close()
    return z


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "golden_multifluid.npz"), **multifluid())
    json.dump(dict(generator="tests/golden/make_golden.py", reference="surajiyer/Fluid-Simulation FluidSystem.cpp (unmodified)",
                   compiler="g++ -std=c++20 -O2 -ffp-contract=off", frames=frames(), bordered_frames=bordered_frames(), mf_frames=mf_frames()),
              open(os.path.join(HERE, "golden_hashes.json"), "w"), indent=1)
    np.savez_compressed(os.path.join(HERE, "golden_fields.npz"), **fields())
    print("wrote", os.listdir(HERE))
