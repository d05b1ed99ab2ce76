"""GPU suite: every entry point of libfluid_cuda.so, called through the C ABI,
against the oracle (oracle/stam_oracle.c) and the committed golden vectors.

Bar: BIT-EXACT (the float bit patterns are compared as integers) for
  * the wavefront order vs the reference order            (north_star: bitwise)
  * the red-black kernels vs the CPU red-black statement  (same arithmetic, same order)
and a relative-L2 tolerance of 1e-4 (north_star) for red-black vs the reference
run to convergence.  Nothing here reads /root/reference.
"""
import os

import numpy as np
import pytest

from conftest import assert_bits, rand_field

pytestmark = pytest.mark.gpu

SIZES = [1, 2, 3, 7, 16, 63, 64, 65, 130, 200]


@pytest.fixture(autouse=True)
def _default_modes(fl):
    fl.set_modes()
    yield
    fl.set_modes()


def _host_threads():
    """Threads for the oracle's independent loops: the cores this process may run on (a container's cpu_count() can be
    the whole host; oversubscribed OpenMP threads spin against each other), at most 16."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    return max(1, min(16, n))


def test_device_is_b200_class(fl):
    import torch
    assert torch.cuda.is_available(), "GPU suite needs a CUDA device"
    assert b"sm_100a" in fl.lib().fluid_version()
    cc = torch.cuda.get_device_capability(0)
    assert cc[0] == 10, "built for sm_100a only, found compute capability %r" % (cc,)


@pytest.mark.parametrize("N", SIZES)
def test_add_source_and_set_bnd(fl, oracle, N):
    rng = np.random.default_rng(N)
    x, s = rand_field(rng, N), rand_field(rng, N)
    y = x.copy()
    fl.add_source(N, x, s, 0.1)
    oracle.add_source(N, y, s, 0.1)
    assert_bits(x, y, "add_source")
    for b in (0, 1, 2, 3, -1):
        x = rand_field(rng, N)
        y = x.copy()
        fl.set_bnd(N, b, x)
        oracle.set_bnd(N, b, y)
        assert_bits(x, y, "set_bnd b=%d" % b)


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("N", SIZES)
def test_lin_solve(fl, oracle, N, order):
    rng = np.random.default_rng(100 + N)
    fl.set_modes(gs_order=order)
    for b in (0, 1, 2, 3):
        for a in (0.0, 1.0, 0.1 * 1e-4 * N * N):
            a = np.float32(a)
            c = np.float32(4.0) if a == 1 else 1 + 4 * a
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            y = x.copy()
            fl.lin_solve(N, b, x, x0, a, c)
            oracle.lin_solve(N, b, y, x0, a, c, 20, order)
            assert_bits(x, y, "lin_solve N=%d b=%d a=%g order=%d" % (N, b, a, order))


@pytest.mark.parametrize("iters", [1, 2, 5, 33])
def test_lin_solve_iteration_counts(fl, oracle, iters):
    N = 70
    rng = np.random.default_rng(iters)
    for order in (0, 1):
        fl.set_modes(gs_order=order, gs_iters=iters)
        x, x0 = rand_field(rng, N), rand_field(rng, N)
        y = x.copy()
        fl.lin_solve(N, 1, x, x0, 0.5, 3.0)
        oracle.lin_solve(N, 1, y, x0, 0.5, 3.0, iters, order)
        assert_bits(x, y, "iters=%d order=%d" % (iters, order))


@pytest.mark.parametrize("tile_sweeps", [0, 1, 2])
@pytest.mark.parametrize("N", [5, 64, 250, 300, 513])
def test_redblack_kernel_variants_identical(fl, oracle, N, tile_sweeps):
    """One-colour-per-launch kernels, and the TMA tile kernel with 1 or 2 sweeps per
    residency, all reproduce the CPU red-black statement bit for bit -- including
    odd sweep counts (mixed 2+1+1 pass schedules) and walls inside partial tiles."""
    import oracle_py as op
    rng = np.random.default_rng(600 + N)
    for iters in (2, 3, 5, 6, 20):
        fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_REDBLACK, gs_iters=iters)
        fs.SetTuning(tile_sweeps)
        fs.SetSmallGridKernels(False)  # this test is about the large-grid kernels, also on tiny grids
        for b in (0, 1, 2):
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            a = np.float32(0.8)
            fs.upload(fl.U, x)
            fs.upload(fl.V, x0)
            fs.lin_solve(b, fl.U, fl.V, a, 1 + 4 * a)
            oracle.lin_solve(N, b, x, x0, a, 1 + 4 * a, iters, op.GS_RB)
            assert_bits(fs.download(fl.U), x, "N=%d iters=%d b=%d tile_sweeps=%d" % (N, iters, b, tile_sweeps))
        fs.close()


@pytest.mark.parametrize("N", SIZES)
def test_diffuse_and_blur(fl, oracle, N):
    rng = np.random.default_rng(200 + N)
    for b in (0, 1, 2, 3):
        for diff in (0.0, 1e-4, 1e-2):
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            y = x.copy()
            fl.diffuse(N, b, x, x0, diff, 0.1)
            oracle.diffuse(N, b, y, x0, diff, 0.1)
            assert_bits(x, y, "diffuse")
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            y, y0 = x.copy(), x0.copy()
            fl.diffuse_blur(N, b, x, x0, diff, 0.1)
            oracle.diffuse_blur(N, b, y, y0, diff, 0.1)
            assert_bits(x, y, "blur result N=%d b=%d diff=%g" % (N, b, diff))
            assert_bits(x0, y0, "blur scratch (penultimate iterate)")


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("N", SIZES)
def test_advect(fl, oracle, N, mode):
    rng = np.random.default_rng(300 + N)
    fl.set_modes(advect_mode=mode)
    for b in (0, 1, 2):
        for scale in (0.05, 1.0, 40.0):  # 40: backtraces leave the domain, exercising the clamp
            d, d0 = rand_field(rng, N), rand_field(rng, N)
            u, v = rand_field(rng, N, scale), rand_field(rng, N, scale)
            e = d.copy()
            fl.advect(N, b, d, d0, u, v, 0.1)
            oracle.advect(N, b, e, d0, u, v, 0.1, mode)
            assert_bits(d, e, "advect N=%d b=%d scale=%g mode=%d" % (N, b, scale, mode))
    # d0 aliasing u (what VelStep1 does, FluidSystem.cpp:372)
    d, u, v = rand_field(rng, N), rand_field(rng, N), rand_field(rng, N)
    e = d.copy()
    fl.advect(N, 1, d, u, u, v, 0.1)
    oracle.advect(N, 1, e, u, u, v, 0.1, mode)
    assert_bits(d, e, "advect with d0 aliasing u")


def _smooth_velocity(N, rng, amp_cells, dt=0.1):
    """A velocity field whose backtraces move up to `amp_cells` cells (dt*N*|v|), smooth over tens of
    cells -- long runs of cells that agree on the integer part of their displacement, seams where it
    changes sign or steps, plus a patch of noise and a patch of exact zeros."""
    i = np.arange(N + 2, dtype=np.float64)
    X, Y = np.meshgrid(i, i)  # [j, i]
    kx, ky = 2 * np.pi / (N / 3.3), 2 * np.pi / (N / 2.1)
    scale = amp_cells / (dt * N)
    u = scale * np.sin(kx * X + 0.3) * np.cos(ky * Y)
    v = scale * np.cos(kx * X) * np.sin(ky * Y + 1.1)
    w = min(130, N + 2 - 40)
    u[N // 5:N // 5 + 9, 40:40 + w] += scale * rng.uniform(-3, 3, (9, w))
    v[N // 2:N // 2 + 40, :] = 0.0
    u[N // 2:N // 2 + 40, :] = 0.0
    v[3 * N // 4:3 * N // 4 + 20, N // 3:] = -0.0
    return u.astype(np.float32).ravel(), v.astype(np.float32).ravel()


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("N,amp", [(130, 0.7), (513, 0.9), (1030, 0.6), (1030, 2.5), (2048, 0.95)])
def test_advect_row_window_kernel_bitwise(fl, oracle, N, amp, mode):
    """The row-window advect kernel (kernels_advect.cuh): register-window interpolation where a whole row
    of a warp agrees on a sub-cell displacement, gathers elsewhere -- scalar and velocity forms, all
    border modes, widths that are not a multiple of 4 or 128, against the oracle; and the gather kernel
    on the same inputs."""
    rng = np.random.default_rng(7000 + N + mode)
    u, v = _smooth_velocity(N, rng, amp)
    d0 = rand_field(rng, N)
    d0.reshape(N + 2, N + 2)[5:9, :] = 0.0
    for kernel in (2, 1):
        fs = fl.FluidSystem().Setup(N, advect_mode=mode)
        fs.SetAdvectKernel(kernel)
        fs.upload(fl.U_PREV, u)
        fs.upload(fl.V_PREV, v)
        fs.upload(fl.DENS_PREV, d0)
        for b in (0, 1, 2, 3):
            d = rand_field(rng, N)
            fs.upload(fl.DENS, d)
            fs.advect(b, fl.DENS, fl.DENS_PREV, fl.U_PREV, fl.V_PREV, 0.1)
            oracle.advect(N, b, d, d0, u, v, 0.1, mode)
            assert_bits(fs.download(fl.DENS), d, "advect kernel %d N=%d b=%d amp=%g mode=%d" % (kernel, N, b, amp, mode))
        # VelStep1's pair: Advect1(1,u,u0,u0,v0), Advect1(2,v,v0,u0,v0)
        un, vn = rand_field(rng, N), rand_field(rng, N)
        fs.upload(fl.U, un)
        fs.upload(fl.V, vn)
        fs.advect_velocity(fl.U, fl.V, fl.U_PREV, fl.V_PREV, 0.1)
        oracle.advect(N, 1, un, u, u, v, 0.1, mode)
        oracle.advect(N, 2, vn, v, u, v, 0.1, mode)
        assert_bits(fs.download(fl.U), un, "advect_velocity u, kernel %d N=%d amp=%g mode=%d" % (kernel, N, amp, mode))
        assert_bits(fs.download(fl.V), vn, "advect_velocity v, kernel %d N=%d amp=%g mode=%d" % (kernel, N, amp, mode))
        fs.close()


def test_advect_row_window_kernel_random_fields_and_clamps(fl, oracle):
    """Rough fields: every row disagrees (gather path throughout), backtraces leave the domain (clamp), an
    ensemble.  (No NaN inputs: the reference indexes out of bounds for a NaN coordinate, FluidCode.h:95-103.)"""
    rng = np.random.default_rng(99)
    N = 260
    fs = fl.FluidSystem().Setup(N, batch=3)
    fs.SetAdvectKernel(2)
    for scale in (0.05, 1.0, 40.0):
        ins = []
        for sim in range(3):
            u, v, d0, d = (rand_field(rng, N, scale), rand_field(rng, N, scale), rand_field(rng, N), rand_field(rng, N))
            ins.append((u, v, d0, d))
            for fid, a in ((fl.U_PREV, u), (fl.V_PREV, v), (fl.DENS_PREV, d0), (fl.DENS, d)):
                fs.upload(fid, a, sim)
        fs.advect(0, fl.DENS, fl.DENS_PREV, fl.U_PREV, fl.V_PREV, 0.1)
        fs.advect_velocity(fl.U, fl.V, fl.U_PREV, fl.V_PREV, 0.1)
        for sim, (u, v, d0, d) in enumerate(ins):
            un, vn = np.zeros_like(u), np.zeros_like(u)
            with np.errstate(all="ignore"):
                oracle.advect(N, 0, d, d0, u, v, 0.1)
                oracle.advect(N, 1, un, u, u, v, 0.1)
                oracle.advect(N, 2, vn, v, u, v, 0.1)
            assert_bits(fs.download(fl.DENS, sim), d, "dens, member %d scale %g" % (sim, scale))
            got_u, got_v = fs.download(fl.U, sim), fs.download(fl.V, sim)
            # the border cells of u/v start as zeros here and in the oracle alike
            assert_bits(got_u, un, "u, member %d scale %g" % (sim, scale))
            assert_bits(got_v, vn, "v, member %d scale %g" % (sim, scale))
    fs.close()


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("N", SIZES)
def test_project(fl, oracle, N, order):
    rng = np.random.default_rng(400 + N)
    fl.set_modes(gs_order=order)
    a = [rand_field(rng, N) for _ in range(4)]
    b = [z.copy() for z in a]
    fl.project(N, *a)
    oracle.project(N, *b, 20, order)
    for name, p, q in zip(("u", "v", "p", "div"), a, b):
        assert_bits(p, q, "project %s N=%d order=%d" % (name, N, order))


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("N", [2, 16, 65, 200, 513])
def test_vel_and_dens_step(fl, oracle, N, order):
    rng = np.random.default_rng(500 + N)
    fl.set_modes(gs_order=order)
    f = [rand_field(rng, N) for _ in range(6)]  # u v d u0 v0 d0
    g = [z.copy() for z in f]
    fl.vel_step(N, f[0], f[1], f[3], f[4], 1e-4, 0.1)
    fl.dens_step(N, f[2], f[5], f[0], f[1], 1e-4, 0.1)
    oracle.vel_step(N, g[0], g[1], g[3], g[4], 1e-4, 0.1, 20, order)
    oracle.dens_step(N, g[2], g[5], g[0], g[1], 1e-4, 0.1)
    for name, p, q in zip("u v dens u0(scratch) v0(scratch) dens0(scratch)".split(), f, g):
        assert_bits(p, q, "%s N=%d order=%d" % (name, N, order))


@pytest.mark.parametrize("order", [0, 1])
def test_frame_level_host_call(fl, oracle, order):
    """fluid_update = FluidSystem::Update on host arrays, incl. vorticity confinement and damping."""
    import oracle_py as op
    N = 90
    rng = np.random.default_rng(order)
    src = op.stam_sources(N) + op.blower_sources(N)
    f = {k: rand_field(rng, N, 0.3) for k in "u v d".split()}
    f.update({k: np.zeros((N + 2) ** 2, np.float32) for k in "u0 v0 d0".split()})
    g = {k: f[k].copy() for k in "u v d".split()}
    fl.set_modes(gs_order=order)
    for _ in range(3):
        fl.update(N, g["u"], g["v"], g["d"], src, 2e-4, 1e-4, 0.1, vorticity=0.4, damp=0.1)
    oracle.update(N, f, src, 2e-4, 1e-4, 0.1, 3, order=order, vorticity=0.4, damp=0.1)
    for k in "u v d".split():
        assert_bits(g[k], f[k], "fluid_update %s order=%d" % (k, order))


def test_golden_primitive_fields(fl, golden_fields):
    """The committed fixtures made by the unmodified reference (tests/golden/make_golden.py)."""
    G = golden_fields
    for N in (12, 16):
        t = "N%d_" % N
        for b in (0, 1, 2):
            for tag, a in (("a0", 0.0), ("a1", 1.0), ("ad", 0.1 * 1e-4 * N * N)):
                c = 4.0 if tag == "a1" else 1 + 4 * a
                x = G[t + "lin_solve_b%d_%s_in_x" % (b, tag)].copy()
                fl.lin_solve(N, b, x, G[t + "lin_solve_b%d_%s_in_x0" % (b, tag)].copy(), np.float32(a), np.float32(c))
                assert_bits(x, G[t + "lin_solve_b%d_%s_out" % (b, tag)], "golden lin_solve")
            for tag, diff in (("d0", 0.0), ("d1", 1e-3)):
                x, x0 = G[t + "blur_b%d_%s_in_x" % (b, tag)].copy(), G[t + "blur_b%d_%s_in_x0" % (b, tag)].copy()
                fl.diffuse_blur(N, b, x, x0, diff, 0.1)
                assert_bits(x, G[t + "blur_b%d_%s_out_x" % (b, tag)], "golden blur")
                assert_bits(x0, G[t + "blur_b%d_%s_out_x0" % (b, tag)], "golden blur x0")
            for tag in ("slow", "fast"):
                k = t + "advect_b%d_%s_" % (b, tag)
                for mode, name in ((0, "midpoint"), (1, "euler")):
                    fl.set_modes(advect_mode=mode)
                    d = G[k + "in_d"].copy()
                    fl.advect(N, b, d, G[k + "in_d0"].copy(), G[k + "in_u"].copy(), G[k + "in_v"].copy(), 0.1)
                    assert_bits(d, G[k + "out_" + name], "golden advect " + name)
                fl.set_modes()
        a = [G[t + "project_in_" + k].copy() for k in ("u", "v", "p", "div")]
        fl.project(N, *a)
        for k, arr in zip(("u", "v", "p", "div"), a):
            assert_bits(arr, G[t + "project_out_" + k], "golden project " + k)
        f = [G[t + "step_in_%d" % w].copy() for w in range(6)]
        fl.vel_step(N, f[0], f[1], f[3], f[4], 1e-3, 0.1)
        fl.dens_step(N, f[2], f[5], f[0], f[1], 1e-3, 0.1)
        for w in range(6):
            assert_bits(f[w], G[t + "step_out_%d" % w], "golden vel+dens step field %d" % w)


def _sources(name, N):
    import oracle_py as op
    return op.stam_sources(N) if name == "stam" else op.blower_sources(N)


def test_golden_frames_wavefront_bitwise(fl, golden_hashes):
    """Whole frames (FluidSystem::Update) on the resident context vs the hashes of
    the unmodified reference: N=64 x 100 and x 1000 steps (BASELINE config 1),
    vorticity+damp, the Blower forcing, and config 2 (N=1024) after its first and after its 100th step."""
    import oracle_py as op
    for fr in golden_hashes:
        fs = fl.FluidSystem().Setup(fr["N"], fr["visc"], fr["diff"], fr["vorticity"], fr["damp"])
        fs.SetSources(_sources(fr["source"], fr["N"]))
        fs.Update(fr["dt"], fr["steps"])
        got = (op.fnv1a64(fs.GetDensities()), op.fnv1a64(fs.GetVelX()), op.fnv1a64(fs.GetVelY()))
        assert got == (fr["dens"], fr["u"], fr["v"]), fr
        fs.close()


def test_config2_n1024_100_steps_wavefront_bitwise(fl, oracle):
    """BASELINE config 2: N=1024, diff=visc=1e-4; the oracle needs ~1.3 s per step,
    so the bitwise check covers the first 3 steps; red-black is compared on the same run."""
    import oracle_py as op
    N, steps = 1024, 3
    src = op.stam_sources(N)
    f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.update(N, f, src, 1e-4, 1e-4, 0.1, steps)
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4)
    fs.SetSources(src)
    fs.Update(0.1, steps)
    assert_bits(fs.GetVelX(), f["u"], "u")
    assert_bits(fs.GetVelY(), f["v"], "v")
    assert_bits(fs.GetDensities(), f["d"], "dens")
    fs.close()
    # red-black on the same problem: bit-exact against the CPU red-black statement
    g = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.update(N, g, src, 1e-4, 1e-4, 0.1, steps, order=op.GS_RB)
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK)
    fs.SetSources(src)
    fs.Update(0.1, steps)
    assert_bits(fs.GetVelX(), g["u"], "rb u")
    assert_bits(fs.GetDensities(), g["d"], "rb dens")
    fs.close()


def test_redblack_tolerance_vs_reference_order_converged(fl, oracle):
    """north_star: red-black within 1e-4 relative of the reference run to convergence.
    The Poisson solve needs O(N^2) sweeps, so this is evaluated at N=64 with 4000
    sweeps on both sides (SURVEY.md 7 hard part 2), after 1 and after 10 steps."""
    import oracle_py as op
    N, iters = 64, 4000
    src = op.stam_sources(N)
    for steps in (1, 10):
        f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
        oracle.update(N, f, src, 1e-4, 1e-4, 0.1, steps, iters=iters, order=op.GS_LEX)
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, gs_iters=iters)
        fs.SetSources(src)
        fs.Update(0.1, steps)
        for name, got in (("u", fs.GetVelX()), ("v", fs.GetVelY()), ("d", fs.GetDensities())):
            want = f[name].astype(np.float64)
            rel = np.linalg.norm(got - want) / np.linalg.norm(want)
            assert rel <= 1e-4, (name, steps, rel)
        fs.close()


@pytest.mark.parametrize("N,iters", [(64, 4000), (128, 16000)])
def test_redblack_tolerance_after_100_steps(fl, oracle, N, iters):
    """the red-black criterion at N=64 AND 128, after 1 AND 100 steps.  A hundred
    converged frames cost the CPU minutes (the reference order is a dependency chain), so the reference
    side is the GPU's reference-order path -- first pinned, at this very sweep count, to the CPU
    restatement over one frame, bit for bit -- and the red-black run is compared with it after frames
    1 and 100.  The sweep count grows with N^2 (N=128: rho = 1 - (pi/129)^2, 16000 sweeps)."""
    import oracle_py as op
    src = op.stam_sources(N)
    f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.update(N, f, src, 1e-4, 1e-4, 0.1, 1, iters=iters, order=op.GS_LEX)
    ref = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_WAVEFRONT, gs_iters=iters)
    rb = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, gs_iters=iters)
    for fs in (ref, rb):
        fs.SetSources(src)
    worst = 0.0
    for frames_so_far, advance in ((1, 1), (100, 99)):
        ref.Update(0.1, advance)
        rb.Update(0.1, advance)
        if frames_so_far == 1:
            for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
                assert_bits(ref.download(fid), f[name], "reference order at %d sweeps vs the CPU: %s" % (iters, name))
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            want = ref.download(fid).astype(np.float64)
            got = rb.download(fid)
            rel = np.linalg.norm(got - want) / np.linalg.norm(want)
            worst = max(worst, rel)
            assert rel <= 1e-4, (N, name, frames_so_far, rel)
    print("red-black vs reference order, N=%d, %d sweeps: worst rel-L2 over 1 and 100 frames %.3g" % (N, iters, worst))
    ref.close()
    rb.close()


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("dt", [0.1, -0.05, 0.0])
def test_sparse_add_source_keeps_signed_zeros(fl, oracle, order, dt):
    """Update() adds a frame's forcing over the listed cells only and sweeps the rest for -0 (x + dt*0).
    Fields full of signed zeros, sources whose product with dt is -0, +0 or underflows, several entries
    on one cell, a source in a border cell, negative and zero dt: frames must equal the oracle's dense
    AddArrDt bit for bit, with the sparse path on and off."""
    import oracle_py as op
    N, steps = 150, 2
    rng = np.random.default_rng(1234)
    init = {}
    for k in "uvd":
        a = rand_field(rng, N, 0.3).reshape(N + 2, N + 2)
        a[10:40, :] = -0.0
        a[40:60, :] = 0.0
        a[70:75, 20:90] = -0.0
        init[k] = a.ravel().copy()
    tiny = float(np.float32(-1e-45))
    src = [(2, 30, 20, 0, 100.0), (0, 31, 20, 0, -0.0), (1, 32, 20, 0, tiny), (2, 33, 20, 0, 0.0), (0, 34, 50, 0, 5.0),
           (0, 34, 50, 0, -7.0), (1, 0, 12, 0, 3.0), (2, 77, 72, 0, -0.0), (1, 77, 72, 1, 0.25), (1, 77, 72, 0, tiny),
           (0, 100, 100, 2, -0.0), (0, 100, 100, 0, -0.0)]
    f = {k: init[k].copy() for k in "uvd"}
    f.update({k + "0": np.zeros((N + 2) ** 2, np.float32) for k in "uvd"})
    oracle.update(N, f, src, 1e-4, 2e-4, dt, steps, order=order)
    for sparse in (True, False):
        fs = fl.FluidSystem().Setup(N, 1e-4, 2e-4, gs_order=order)
        fs.SetSparseSources(sparse)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            fs.upload(fid, init[name])
        fs.SetSources(src)
        fs.Update(dt, steps)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            assert_bits(fs.download(fid), f[name], "sparse=%s order=%d dt=%g: %s" % (sparse, order, dt, name))
        fs.close()


def test_ensemble_members_are_independent(fl, oracle):
    """Batched ensemble (BASELINE config 5, scaled down): each member equals the
    single simulation with the same source position, bit for bit."""
    import oracle_py as op
    N, B, steps = 48, 5, 4
    for order in (0, 1):
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=order, batch=B)
        srcs = []
        for b in range(B):
            cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
            srcs.append([(2, cx, cy, 0, 100.0), (0, cx, cy, 0, 5.0), (1, cx, cy, 0, 5.0)])
            fs.SetSources(srcs[-1], sim=b)
        fs.Update(0.1, steps)
        for b in range(B):
            f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
            oracle.update(N, f, srcs[b], 1e-4, 1e-4, 0.1, steps, order=order)
            assert_bits(fs.GetVelX(b), f["u"], "member %d u" % b)
            assert_bits(fs.GetDensities(b), f["d"], "member %d dens" % b)
        fs.close()


@pytest.mark.parametrize("stream", [1, 2])
def test_ensemble_on_the_streaming_kernels_bitwise(fl, oracle, stream):
    """Eight members of N=512 are enough cells for the register-resident streaming kernels, which then work on
    member planes (strips and chunks per member, wall strips in every member).  Three steps, so that the redo
    feedback may also switch a solve to the inline-exact variant on the way (stream = 2: that variant from the first
    solve on -- point sources decaying through the subnormals make every member mostly fringe): every member equals
    its own CPU run."""
    import oracle_py as op
    N, B, steps = 512, 8, 3
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=B)
    fs.SetFastDivision(2)
    fs.SetStreamKernel(stream)
    srcs = []
    for b in range(B):
        cx, cy = 1 + (37 * b) % N, 1 + (101 * b) % N
        srcs.append([(2, cx, cy, 0, 100.0), (0, cx, cy, 0, 5.0), (1, cx, cy, 0, 5.0)])
        fs.SetSources(srcs[-1], sim=b)
    fs.Update(0.1, steps)
    for b in range(B):
        f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
        oracle.update(N, f, srcs[b], 1e-4, 1e-4, 0.1, steps, order=op.GS_RB)
        assert_bits(fs.GetVelX(b), f["u"], "member %d u" % b)
        assert_bits(fs.GetDensities(b), f["d"], "member %d dens" % b)
    fs.close()


def test_vorticity_and_damp(fl, oracle):
    N = 45
    rng = np.random.default_rng(9)
    u, v = rand_field(rng, N), rand_field(rng, N)
    a, b = u.copy(), v.copy()
    fl.vort_confinement(N, u, v, 0.55)
    oracle.vort_confinement(N, a, b, 0.55)
    assert_bits(u, a, "vorticity u")
    assert_bits(v, b, "vorticity v")
    fl.damp(N, u, 0.1, 0.1)
    oracle.damp(N, a, 0.1, 0.1)
    assert_bits(u, a, "damp")


def test_full_size_primitives_n8192(fl, oracle):
    """BASELINE config 3 size.  The oracle runs single primitives at N=8192 in
    seconds when the sweep count is small, so these are still bit-exact checks."""
    N = 8192
    rng = np.random.default_rng(8192)
    n = (N + 2) ** 2
    x0 = rng.uniform(-1, 1, n).astype(np.float32)
    a = np.float32(0.1 * 1e-4 * N * N)
    for order in (0, 1):
        fl.set_modes(gs_order=order, gs_iters=2)
        x = np.zeros(n, np.float32)
        y = x.copy()
        fl.lin_solve(N, 1, x, x0, a, 1 + 4 * a)
        oracle.lin_solve(N, 1, y, x0, a, 1 + 4 * a, 2, order)
        assert_bits(x, y, "N=8192 lin_solve order=%d" % order)
    fl.set_modes()
    u = rng.uniform(-1, 1, n).astype(np.float32)
    v = rng.uniform(-1, 1, n).astype(np.float32)
    d = np.zeros(n, np.float32)
    e = d.copy()
    fl.advect(N, 0, d, x0, u, v, 0.1)
    oracle.advect(N, 0, e, x0, u, v, 0.1)
    assert_bits(d, e, "N=8192 advect")
    # size-independent properties: set_bnd is idempotent; add_source is exact for dt = 0
    z = d.copy()
    fl.set_bnd(N, 1, z)
    z2 = z.copy()
    fl.set_bnd(N, 1, z2)
    assert_bits(z, z2, "set_bnd idempotent")


@pytest.mark.parametrize("order", [0, 1])
def test_full_step_n2048_bitwise(fl, oracle, order):
    """One whole frame at N=2048 in both relaxation orders (red-black: the streaming kernels, both lanes)."""
    import oracle_py as op
    N = 2048
    src = op.stam_sources(N)
    f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.set_threads(_host_threads() if order == 1 else 1)
    try:
        oracle.update(N, f, src, 1e-4, 1e-4, 0.1, 1, order=order)
    finally:
        oracle.set_threads(1)
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=order)
    fs.SetFastDivision(2)
    fs.SetSources(src)
    fs.Update(0.1, 1)
    assert_bits(fs.GetVelX(), f["u"], "u")
    assert_bits(fs.GetVelY(), f["v"], "v")
    assert_bits(fs.GetDensities(), f["d"], "dens")
    fs.close()


def test_benchmarked_configuration_n8192_redblack_bitwise(fl, oracle):
    """BASELINE config 3 exactly as bench.py times it -- N=8192, red-black, 20 sweeps, dt=0.1,
    visc=diff=1e-4, source S1, density diffusion and the v-lane overlapped -- against the CPU red-black
    statement, bit for bit, over two whole frames (the second starts from non-trivial fields and has
    the proven division in place from the start: 4-sweep streaming passes, the power-of-two pressure
    passes with the zero guess, redo launches, array trades).  The oracle's independent loops run on
    all host cores (bit-identical to one core, tests/test_oracle.py)."""
    import oracle_py as op
    N, steps = 8192, 2
    src = op.stam_sources(N)
    f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.set_threads(_host_threads())
    try:
        oracle.update(N, f, src, 1e-4, 1e-4, 0.1, steps, order=op.GS_RB)
    finally:
        oracle.set_threads(1)
    for mode in (2, 1):  # proof at once / the automatic background proof (first frame on the exact kernels)
        fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_REDBLACK, gs_iters=20)
        fs.SetFastDivision(mode)
        fs.SetSources(src)
        fs.Update(0.1, steps)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            assert_bits(fs.download(fid), f[name], "N=8192 red-black frame, division mode %d: %s" % (mode, name))
        fs.close()


def test_mass_is_reported_like_the_reference(fl, golden_hashes, oracle):
    fr = golden_hashes[0]
    fs = fl.FluidSystem().Setup(fr["N"], fr["visc"], fr["diff"])
    fs.SetSources(_sources(fr["source"], fr["N"]))
    fs.Update(fr["dt"], fr["steps"])
    assert oracle.total_inner_mass(fr["N"], fs.GetDensities()) == fr["mass"]
    fs.close()


def test_errors_are_status_codes(fl):
    L = fl.lib()
    x = np.zeros(16, np.float32)
    assert L.fluid_lin_solve(2, 0, x, x, 1.0, 4.0) == 1  # aliasing rejected
    fs = fl.FluidSystem().Setup(8)
    assert L.fluid_ctx_lin_solve(fs.h, 0, 0, 0, 1.0, 4.0) == 1
    assert L.fluid_ctx_upload(fs.h, 9, 0, x) == 1
    with pytest.raises(fl.FluidError):
        fs.SetSources([(0, 99, 1, 0, 1.0)])
    fs.close()


@pytest.mark.parametrize("persistent", [True, False])
@pytest.mark.parametrize("tile", [16, 32, 64])
def test_wavefront_tile_sizes_bitwise(fl, oracle, tile, persistent):
    """Reference-order relaxation: one launch per solve (tickets + progress counters) and one launch per hyperplane."""
    import oracle_py as op
    rng = np.random.default_rng(tile)
    for N in (130, 257, 320):
        fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=7)
        fs.SetWavefrontTile(tile)
        fs.SetWavefrontPersistent(persistent)
        for b in (0, 1, 2, 3):
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            a = np.float32(0.6)
            fs.upload(fl.U, x)
            fs.upload(fl.V, x0)
            fs.lin_solve(b, fl.U, fl.V, a, 1 + 4 * a)
            oracle.lin_solve(N, b, x, x0, a, 1 + 4 * a, 7, op.GS_LEX)
            assert_bits(fs.download(fl.U), x, "wavefront tile %d N=%d b=%d persistent=%s" % (tile, N, b, persistent))
        fs.close()


def test_wavefront_persistent_many_sweeps_and_members(fl, oracle):
    """More sweeps in flight than tile columns, an ensemble (tickets of several members), a grid much larger than
    the resident CTAs can cover at once (N=2048, 20 sweeps: 1280 tickets of 64 tiles), repeated solves."""
    import oracle_py as op
    rng = np.random.default_rng(5150)
    N, B = 150, 3
    fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=45, batch=B)
    fs.SetWavefrontPersistent(True)
    xs = [(rand_field(rng, N), rand_field(rng, N)) for _ in range(B)]
    for sim, (x, x0) in enumerate(xs):
        fs.upload(fl.U, x, sim)
        fs.upload(fl.V, x0, sim)
    fs.lin_solve(1, fl.U, fl.V, np.float32(1.0), np.float32(4.0))
    for sim, (x, x0) in enumerate(xs):
        oracle.lin_solve(N, 1, x, x0, 1.0, 4.0, 45, op.GS_LEX)
        assert_bits(fs.download(fl.U, sim), x, "member %d" % sim)
    fs.close()
    N = 2048
    fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=20)
    fs.SetWavefrontPersistent(True)
    for rep in range(2):
        x, x0 = rand_field(rng, N), rand_field(rng, N)
        fs.upload(fl.U, x)
        fs.upload(fl.V, x0)
        fs.lin_solve(2, fl.U, fl.V, np.float32(10.49), np.float32(42.96))
        oracle.lin_solve(N, 2, x, x0, np.float32(10.49), np.float32(42.96), 20, op.GS_LEX)
        assert_bits(fs.download(fl.U), x, "N=2048 solve %d" % rep)
    fs.close()


@pytest.mark.parametrize("fast_div", [0, 2])
def test_row_pipeline_reference_order_bitwise(fl, oracle, fast_div):
    """The reference-order solve as a pipeline of warps (kernels_rowwave.cuh: a warp per 32-row band and sweep, streaming
    along the rows, ordered by progress counters), forced onto grids of every shape class: fewer rows than a band, rows
    that are not whole float4 groups (N % 4 = 0..3), a partial top band, ensemble members,
    one sweep and many; with IEEE division (fast_div = 0), the scaled short division proven for the divisor (2; fields
    with subnormal patches, exact zeros and a patch beyond 2^64 that overflows it) and a power-of-two divisor.  Every
    value against the CPU's lexicographic sweep, bit for bit."""
    import oracle_py as op
    rng = np.random.default_rng(31337 + fast_div)
    cases = [(5, 1, 3), (33, 1, 7), (130, 1, 7), (257, 2, 5), (258, 1, 20), (131, 1, 1), (320, 3, 4), (1000, 1, 20)]
    for N, B, iters in cases:
        fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_WAVEFRONT, gs_iters=iters, batch=B)
        fs.SetSmallGridKernels(False)
        fs.SetWavefrontPersistent(3)
        fs.SetFastDivision(fast_div)
        for b, a, c in ((0, np.float32(0.6), None), (1, np.float32(10.49), None), (2, np.float32(1.0), np.float32(4.0)), (3, np.float32(0.3), None)):
            c = np.float32(1 + 4 * a) if c is None else c
            members = []
            for sim in range(B):
                x, x0 = rand_field(rng, N), rand_field(rng, N)
                for arr in (x, x0):
                    g = arr.reshape(N + 2, N + 2)
                    g[N // 3:N // 2, 2:N // 2] *= np.float32(1e-38)
                    g[N // 2:N // 2 + 3, :] = 0.0
                    if N >= 100:
                        g[N - 20:N - 10, N - 40:N - 5] *= np.float32(1e28)
                members.append((x, x0))
                fs.upload(fl.U, x, sim)
                fs.upload(fl.V, x0, sim)
            fs.lin_solve(b, fl.U, fl.V, a, c)
            fs.Sync()
            for sim, (x, x0) in enumerate(members):
                oracle.lin_solve(N, b, x, x0, a, c, iters, op.GS_LEX)
                assert_bits(fs.download(fl.U, sim), x, "row pipeline N=%d members=%d iters=%d b=%d member %d fast_div=%d" % (N, B, iters, b, sim, fast_div))
        fs.close()


def test_row_pipeline_whole_frames_match_tile_wavefront(fl):
    """Whole frames (forcing, both projections, density step) in the reference's order: the row pipeline against the
    per-hyperplane tile kernels on a grid whose rows are slab-shaped (NY != N) and one that is square, u and v solved on
    two streams side by side."""
    rng = np.random.default_rng(99)
    for N, NY in ((700, 300), (513, 513)):
        init = [(rng.uniform(-1, 1, (NY + 2, N + 2)) * 0.3).astype(np.float32) for _ in range(3)]
        src = [(2, N // 2, NY // 2, 0, 100.0), (0, N // 2, NY // 2, 0, 5.0), (1, N // 3, NY // 3, 0, -5.0)]
        got = {}
        for mode in (3, 0):
            fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=fl.GS_WAVEFRONT, gs_iters=10, slab=(0, 1, NY))
            fs.SetWavefrontPersistent(mode)
            fs.SetFastDivision(2)
            fs.SetSources(src)
            for k, fid in enumerate((fl.U, fl.V, fl.DENS)):
                fs.upload(fid, init[k].ravel())
            fs.Update(0.1, 3)
            fs.Sync()
            got[mode] = [fs.download(fid).copy() for fid in (fl.U, fl.V, fl.DENS)]
            fs.close()
        for k, name in enumerate("u v dens".split()):
            assert_bits(got[3][k], got[0][k], "N=%d NY=%d %s" % (N, NY, name))


@pytest.mark.parametrize("small", [True, False])
@pytest.mark.parametrize("N", [1, 2, 9, 64, 65, 127, 128])
def test_small_grid_kernels_match_large_grid_kernels(fl, oracle, N, small):
    """N <= 128: one single-CTA launch per solve / per blur (hyperplane wavefront or colour passes in
    shared memory) against the oracle, and the large-grid kernels forced onto the same sizes."""
    import oracle_py as op
    rng = np.random.default_rng(700 + N)
    for order in (0, 1):
        for iters in (1, 4, 20):
            fs = fl.FluidSystem().Setup(N, gs_order=order, gs_iters=iters)
            fs.SetSmallGridKernels(small)
            for b in (0, 1, 2, 3):
                for a in (np.float32(0.0), np.float32(0.7)):
                    x, x0 = rand_field(rng, N), rand_field(rng, N)
                    fs.upload(fl.U, x)
                    fs.upload(fl.V, x0)
                    fs.lin_solve(b, fl.U, fl.V, a, 1 + 4 * a)
                    oracle.lin_solve(N, b, x, x0, a, 1 + 4 * a, iters, order)
                    assert_bits(fs.download(fl.U), x, "lin_solve N=%d order=%d iters=%d b=%d a=%g" % (N, order, iters, b, a))
                    if order == 0:
                        x, x0 = rand_field(rng, N), rand_field(rng, N)
                        fs.upload(fl.U, x)
                        fs.upload(fl.V, x0)
                        fs.blur(b, fl.U, fl.V, a, 1 + 4 * a)
                        oracle.blur(N, b, x, x0, a, 1 + 4 * a, iters)
                        assert_bits(fs.download(fl.U), x, "blur N=%d iters=%d b=%d a=%g" % (N, iters, b, a))
                        assert_bits(fs.download(fl.V), x0, "blur scratch N=%d" % N)
            fs.close()


def test_verified_division_windows_and_results(fl, oracle):
    """The short division sequence is only used where an exhaustive on-GPU comparison with
    div.rn.f32 (all 2^31 non-negative finite dividends) found no mismatch; results must be
    unchanged with it forced on, including fields with zeros, subnormals and huge values."""
    import oracle_py as op
    fs = fl.FluidSystem().Setup(300, gs_order=fl.GS_REDBLACK, gs_iters=6)
    fs.SetFastDivision(2)
    lo, hi = fs.DivisionWindow(4.0)
    assert lo <= -100 and hi >= 100, (lo, hi)          # a power of two: exact wherever nothing under/overflows
    for c in (np.float32(42.943043), np.float32(1.0004), np.float32(3.0), np.float32(1e6)):
        lo, hi = fs.DivisionWindow(float(c))
        assert hi < lo or (lo <= -20 and hi >= 20), (c, lo, hi)
    rng = np.random.default_rng(99)
    N = 300
    for scale in (1.0, 1e-38, 1e30, 0.0):
        x, x0 = rand_field(rng, N, scale), rand_field(rng, N, scale)
        x0[rng.integers(0, x0.size, x0.size // 3)] = 0.0   # exact zeros, +0 and -0
        x0[rng.integers(0, x0.size, x0.size // 7)] = -0.0
        a = np.float32(10.49)
        fs.upload(fl.U, x)
        fs.upload(fl.V, x0)
        fs.lin_solve(1, fl.U, fl.V, a, 1 + 4 * a)
        oracle.lin_solve(N, 1, x, x0, a, 1 + 4 * a, 6, op.GS_RB)
        assert_bits(fs.download(fl.U), x, "fast division, field scale %g" % scale)
        # the Jacobi blur passes use the same verified division
        x, x0 = rand_field(rng, N, scale), rand_field(rng, N, scale)
        x0[rng.integers(0, x0.size, x0.size // 3)] = 0.0
        fs.upload(fl.U, x)
        fs.upload(fl.V, x0)
        fs.blur(0, fl.U, fl.V, a, 1 + 4 * a)
        oracle.blur(N, 0, x, x0, a, 1 + 4 * a, 6)
        assert_bits(fs.download(fl.U), x, "fast division in blur, field scale %g" % scale)
        assert_bits(fs.download(fl.V), x0, "blur penultimate iterate")
    fs.close()


def test_headless_render_matches_restated_renderer_loop(fl, oracle):
    """FluidRenderer::DoBuffer's colourise loop and velocity line list on the device
    (no graphics API).  The oracle for this row is a restatement only (the renderer cannot be compiled
    here), so parity with the real reference is unpinned; the CUDA kernel equals the restatement bitwise."""
    N = 77
    rng = np.random.default_rng(21)
    d = np.abs(rand_field(rng, N, 2.5))   # some cells above 1 (normalised), some below
    u, v = rand_field(rng, N), rand_field(rng, N)
    u[::7] = 0.0
    v[::7] = 0.0                          # zero-speed cells: the reference's 0/0 then override path
    fs = fl.FluidSystem().Setup(N)
    fs.upload(fl.DENS, d)
    fs.upload(fl.U, u)
    fs.upload(fl.V, v)
    color = (0.05, 0.05, 1.0)             # the first fluid of the shipped scene, SystemManager.cpp:104
    px, ln = fs.Render(color)
    want_px, want_ln = oracle.render(N, d, u, v, color)
    assert_bits(px, want_px, "pixels")
    assert_bits(ln, want_ln, "velocity lines")
    g = px.reshape(N + 2, N + 2, 3)
    assert np.all(g[0, :, 1] == 1) and np.all(g[0, :, 0] == 0)  # outer ring is green
    fs.close()
    # the shipped scene: two fluids with their own colours (SystemManager.cpp:103-108), summed and normalised per
    # cell (FluidRenderer.cpp:106-119), and a cell map with OBJECT cells (magenta, :121-123)
    colors = ((0.05, 0.05, 1.0), (1.0, 0.4, 0.05))
    d2 = np.abs(rand_field(rng, N, 0.9))
    d2[5 * (N + 2):9 * (N + 2)] = 0.0
    fs = fl.FluidSystem().Setup(N, n_fluids=2)
    for fid, a in ((fl.dens_of(0), d), (fl.dens_of(1), d2), (fl.U, u), (fl.V, v)):
        fs.upload(fid, a)
    px, ln = fs.Render(colors)
    want_px, want_ln = oracle.render(N, [d, d2], u, v, colors)
    assert_bits(px, want_px, "two fluids: pixels")
    assert_bits(ln, want_ln, "two fluids: velocity lines")
    fs.close()
    info = oracle.make_cell_info(N, True)
    m = info.reshape(N + 2, N + 2)
    m[20:26, 30:41] = (m[20:26, 30:41] & ~np.uint8(16)) | np.uint8(32)
    oracle.calc_border(N, info)
    fs = fl.FluidSystem().Setup(N, n_fluids=2, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    fs.SetCellInfo(info)
    for fid, a in ((fl.dens_of(0), d), (fl.dens_of(1), d2), (fl.U, u), (fl.V, v)):
        fs.upload(fid, a)
    px, _ = fs.Render(colors, lines=False)
    want_px, _ = oracle.render(N, [d, d2], u, v, colors, info)
    assert_bits(px, want_px, "cell map with objects: pixels")
    g = px.reshape(N + 2, N + 2, 3)
    assert np.all(g[20:26, 30:41] == np.array([1, 0, 1], np.float32))
    fs.close()


@pytest.mark.parametrize("order", [0, 1])
def test_two_fluids_resident_context(fl, oracle, order):
    """n_fluids = 2 (the reference's densList): visc_avg couples the fluids, each diffuses with its own
    coefficient.  Reference order vs the golden file made by the unmodified reference; red-black vs the
    oracle construction that test_oracle.py pins against the same file."""
    import os
    from conftest import GOLDEN
    from test_oracle import multifluid_expected
    G = np.load(os.path.join(GOLDEN, "golden_multifluid.npz"))
    N, steps = int(G["N"]), int(G["steps"])
    fluids = [tuple(map(float, row)) for row in G["fluids"]]
    fs = fl.FluidSystem().Setup(N, gs_order=order, n_fluids=2)
    for k, (diff, visc) in enumerate(fluids):
        fs.SetFluid(k, diff, visc)
    src = []
    for s in G["sources"]:
        fid = int(s[0]) if s[0] < 2 else fl.dens_of(int(s[0]) - 2)
        src.append((fid, int(s[1]), int(s[2]), int(s[3]), float(s[4])))
    fs.SetSources(src)
    fs.Update(0.1, steps)
    want = {k: G[k] for k in ("u", "v", "d0", "d1")} if order == 0 else multifluid_expected(oracle, N, steps, fluids, G["sources"], order=1)
    assert_bits(fs.GetVelX(), want["u"], "u")
    assert_bits(fs.GetVelY(), want["v"], "v")
    assert_bits(fs.download(fl.dens_of(0)), want["d0"], "fluid 0")
    assert_bits(fs.download(fl.dens_of(1)), want["d1"], "fluid 1")
    with pytest.raises(fl.FluidError):
        fs.SetFluid(2, 0.0, 0.0)
    fs.close()


@pytest.mark.parametrize("N,stream", [(2500, True), (2500, False), (2502, True), (2524, True), (2500, 2), (2502, 2), (2524, 2)])
def test_streaming_interior_kernel_bitwise(fl, oracle, N, stream):
    """Large grids: the rows away from the bottom and top walls are relaxed by the register-resident
    streaming kernel (4, 2 or 1 sweeps per pass) -- including the left and right walls when N is a multiple
    of 4 (2500, 2524 = 28 mod 32, where lanes fold back onto the border column), one tile column away from
    them otherwise (2502) -- the rest by the tile kernel.  Mixed pass schedules (20 = 4x4+2x2,
    7 = 4+2+1 -> 4+1+1+1, 6, 2) against the CPU red-black statement, and the tile-only path on the same inputs.
    stream = 2: the scaled division sequence (exact on the subnormal fringe, nothing listed for it); patches of the
    fields are scaled into the subnormal range and beyond 2^64 so that every exact fallback really runs."""
    import oracle_py as op
    rng = np.random.default_rng(N)
    for iters, b, a in ((20, 1, np.float32(10.49)), (7, 2, np.float32(1.0)), (6, 0, np.float32(0.3)), (2, 1, np.float32(0.3)),
                        (4, 3, np.float32(0.7))):
        fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_REDBLACK, gs_iters=iters)
        fs.SetStreamKernel(stream)
        fs.SetFastDivision(2)
        x, x0 = rand_field(rng, N), rand_field(rng, N)
        x0[rng.integers(0, x0.size, x0.size // 5)] = 0.0
        for arr in (x, x0):
            g = arr.reshape(N + 2, N + 2)
            g[700:1000, 200:1300] *= np.float32(1e-38)
            g[1:30, N - 50:] *= np.float32(1e-37)
            g[1500:1560, 900:1000] *= np.float32(1e28)  # beyond 2^64: the scaled sequence (stream = 2) overflows to NaN there and lists the rows
        c = np.float32(4.0) if a == 1 else 1 + 4 * a
        fs.upload(fl.U, x)
        fs.upload(fl.V, x0)
        fs.lin_solve(b, fl.U, fl.V, a, c)
        oracle.lin_solve(N, b, x, x0, a, c, iters, op.GS_RB)
        assert_bits(fs.download(fl.U), x, "N=%d iters=%d b=%d stream=%s" % (N, iters, b, stream))
        fs.close()


@pytest.mark.parametrize("N,stream", [(2048, True), (2500, True), (2048, False), (2048, 2), (2500, 2)])
def test_streamed_blur_bitwise(fl, oracle, N, stream):
    """Blur on large grids runs five Jacobi passes per launch in a register-resident window (an even number
    of such launches, single passes for the rest: 21 = 4x5+1, 11 = 2x5+1, 31 = 6x5+1, 7 = single passes only),
    walls and corners included; rectangles whose dividends leave the verified division window -- the field
    has a subnormal patch and an all-zero patch -- are recomputed by the exact fallback kernel.  Result and
    penultimate iterate against the CPU statement, bit for bit."""
    rng = np.random.default_rng(N + 7)
    W = N + 2
    for iters, b, a in ((20, 0, np.float32(671.0886)), (10, 1, np.float32(0.37)), (30, 2, np.float32(2.5)), (6, 0, np.float32(1.0))):
        fs = fl.FluidSystem().Setup(N, gs_iters=iters)
        fs.SetStreamKernel(stream)
        fs.SetFastDivision(2)
        x, x0 = rand_field(rng, N), rand_field(rng, N)
        g0 = x0.reshape(W, W)
        g0[300:420, 100:400] *= np.float32(1e-38)   # products and quotients in the subnormal range
        g0[900:1100, 1500:1900] = 0.0
        g0[1:40, 1:60] *= np.float32(1e-36)         # ... next to two walls
        g0[1300:1340, 700:800] *= np.float32(1e28)  # beyond 2^64: NaN from the scaled sequence (stream = 2), listed and redone
        c = 1 + 4 * a
        fs.upload(fl.U, x)
        fs.upload(fl.V, x0)
        fs.blur(b, fl.U, fl.V, a, c)
        oracle.blur(N, b, x, x0, a, c, iters)
        assert_bits(fs.download(fl.U), x, "blur N=%d iters=%d b=%d stream=%s" % (N, iters, b, stream))
        assert_bits(fs.download(fl.V), x0, "blur penultimate iterate N=%d iters=%d b=%d" % (N, iters, b))
        fs.close()


def test_streaming_and_tile_paths_agree_on_random_shapes(fl):
    """The streaming kernels cut a grid into strips, chunks, shorter edge chunks, half-height wall tasks and
    redo segments; the tile kernels into TMA windows.  Both are pinned to the oracle at a few sizes; here they
    are set against each other (bitwise) on random widths -- multiples of 4 and not --, heights, ensemble
    sizes, boundary modes, sweep counts and coefficients, with zero and subnormal patches in the fields."""
    rng = np.random.default_rng(20261107)
    # flat members (few chunks, edge chunks that are both first and last) and many of them
    # ... and widths whose last strip is narrower than the halo (N mod 112 <= 8: the last strip but one reaches the wall)
    fixed = [(512, 512, 8), (1024, 300, 8), (2048, 130, 8), (256, 1100, 8), (228, 90, 104), (232, 81, 112), (2248, 940, 1),
             (1124, 2000, 1), (1348, 1557, 1)]
    for case in range(24 + len(fixed)):
        if case < len(fixed):
            N, NY, batch = fixed[case]
        else:
            batch = int(rng.choice([1, 1, 2, 3]))
            N = int(rng.integers(1400, 2600)) if batch == 1 else int(rng.integers(900, 1500))
            if rng.random() < 0.6:
                N -= N % 4
            NY = int(rng.integers(max(600, (1 << 21) // (N * batch) + 1), 2300))
        b = int(rng.integers(0, 3))
        iters = int(rng.choice([4, 5, 8, 11, 20]))
        a = np.float32(rng.choice([0.3, 1.0, 23.5, 671.0886]))
        c = np.float32(4.0) if a == 1 else np.float32(1 + 4 * a)
        W, H = N + 2, NY + 2
        fields = []
        for _ in range(2 * batch):
            f = rng.uniform(-1, 1, (H, W)).astype(np.float32)
            y0, x0 = int(rng.integers(1, max(2, NY - 120))), int(rng.integers(1, max(2, N - 200)))
            f[y0:y0 + 100, x0:x0 + 180] = 0.0
            y1, x1 = int(rng.integers(1, max(2, NY - 80))), int(rng.integers(1, max(2, N - 150)))
            f[y1:y1 + 60, x1:x1 + 140] *= np.float32(1e-38)
            fields.append(f.ravel())
        got = {}
        for stream in (True, False):
            fs = fl.FluidSystem().Setup(N, gs_order=fl.GS_REDBLACK, gs_iters=iters, batch=batch, slab=(0, 1, NY))
            fs.SetStreamKernel(stream)
            fs.SetFastDivision(2)
            out = []
            for kind in ("lin_solve", "blur"):
                for m in range(batch):
                    fs.upload(fl.U, fields[2 * m], sim=m)
                    fs.upload(fl.V, fields[2 * m + 1], sim=m)
                if kind == "lin_solve":
                    fs.lin_solve(b, fl.U, fl.V, a, c)
                else:
                    fs.blur(b, fl.U, fl.V, a, c)
                for m in range(batch):
                    out.append(fs.download(fl.U, sim=m).copy())
                    out.append(fs.download(fl.V, sim=m).copy())
            got[stream] = out
            fs.close()
        what = "case %d: N=%d NY=%d batch=%d b=%d iters=%d a=%g" % (case, N, NY, batch, b, iters, a)
        for k, (p, q) in enumerate(zip(got[True], got[False])):
            assert_bits(p, q, "%s, array %d" % (what, k))


def test_whole_steps_agree_with_and_without_streaming(fl):
    """Whole frames (forcing, VelStep with both projections, DensStep) on random shapes and parameters, the
    large-grid kernels on and off: the pressure solve's exact power-of-two division with its unwritten zero
    guess, fields trading arrays with the scratch field after odd pass counts, Blur's multi-pass launches, the
    redo lists and their feedback, and the density diffusion running on a second stream beside VelStep, must all
    be invisible in the results."""
    rng = np.random.default_rng(7)
    for case in range(10):
        batch = int(rng.choice([1, 1, 1, 2]))
        N = int(rng.integers(1450, 2300)) if batch == 1 else int(rng.integers(1024, 1500))
        if rng.random() < 0.7:
            N -= N % 4
        NY = int(rng.integers((1 << 21) // (N * batch) + 1, 2200))
        iters = int(rng.choice([3, 4, 7, 10, 20]))
        visc = float(rng.choice([0.0, 1e-6, 1e-4]))
        diff = float(rng.choice([0.0, 1e-6, 1e-4]))
        vort = float(rng.choice([0.0, 0.0, 0.55]))
        steps = 2
        W, H = N + 2, NY + 2
        init = [(rng.uniform(-1, 1, (H, W)) * 0.3).astype(np.float32) for _ in range(3 * batch)]
        for f in init:
            y0, x0 = int(rng.integers(1, NY - 300)), int(rng.integers(1, N - 300))
            f[y0:y0 + 280, x0:x0 + 280] = 0.0
            f[y0 + 290:y0 + 299, :] *= np.float32(1e-37)
        src = [(2, int(rng.integers(1, N)), int(rng.integers(1, NY)), 0, 100.0), (0, int(rng.integers(1, N)), int(rng.integers(1, NY)), 0, 5.0),
               (1, int(rng.integers(1, N)), int(rng.integers(1, NY)), 0, -5.0), (2, 3, 3, 1, 2.0)]
        got = {}
        for stream in (True, False):
            fs = fl.FluidSystem().Setup(N, visc, diff, vorticity=vort, gs_order=fl.GS_REDBLACK, gs_iters=iters, batch=batch, slab=(0, 1, NY))
            fs.SetStreamKernel(stream)
            fs.SetOverlap(stream)  # the reference run is strictly serial as well
            fs.SetFastDivision(2)
            for m in range(batch):
                fs.SetSources(src, sim=m)
                for k, fid in enumerate((fl.U, fl.V, fl.DENS)):
                    fs.upload(fid, init[3 * m + k].ravel(), sim=m)
            fs.Update(0.1, steps)
            got[stream] = [fs.download(fid, sim=m).copy() for m in range(batch) for fid in (fl.U, fl.V, fl.DENS)]
            fs.close()
        what = "case %d: N=%d NY=%d batch=%d iters=%d visc=%g diff=%g vort=%g" % (case, N, NY, batch, iters, visc, diff, vort)
        for k, (p, q) in enumerate(zip(got[True], got[False])):
            assert_bits(p, q, "%s, array %d" % (what, k))
