"""CPU suite, part 1: the oracle (oracle/stam_oracle.c) against the golden
vectors generated from the unmodified reference, and against the live reference
build when it is present.  Bit-exact everywhere (integer compare of the float
bit patterns)."""
import numpy as np
import pytest

from conftest import assert_bits, rand_field


def _sources(name, N):
    import oracle_py as op
    return op.stam_sources(N) if name == "stam" else op.blower_sources(N)


def test_oracle_frames_match_golden_hashes(oracle, golden_hashes):
    import oracle_py as op
    for fr in golden_hashes:
        if fr["N"] * fr["N"] * fr["steps"] > 64 * 64 * 1000 + 1:
            continue  # N=1024 (1.3 s) is covered in the GPU suite's oracle leg; keep the CPU suite short
        N = fr["N"]
        f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
        oracle.update(N, f, _sources(fr["source"], N), fr["visc"], fr["diff"], fr["dt"], fr["steps"],
                      vorticity=fr["vorticity"], damp=fr["damp"])
        got = (op.fnv1a64(f["d"]), op.fnv1a64(f["u"]), op.fnv1a64(f["v"]))
        assert got == (fr["dens"], fr["u"], fr["v"]), fr
        assert oracle.total_inner_mass(N, f["d"]) == fr["mass"]


def test_oracle_primitives_match_golden_fields(oracle, golden_fields):
    import oracle_py as op
    G = golden_fields
    for N in (12, 16):
        t = "N%d_" % N
        x = G[t + "add_source_in_x"].copy()
        oracle.add_source(N, x, G[t + "add_source_in_s"], 0.1)
        assert_bits(x, G[t + "add_source_out"], "add_source")
        for b in (0, 1, 2, 3):
            x = G[t + "set_bnd_b%d_in" % b].copy()
            oracle.set_bnd(N, b, x)
            assert_bits(x, G[t + "set_bnd_b%d_out" % b], "set_bnd b=%d" % b)
        for b in (0, 1, 2):
            for tag, a in (("a0", 0.0), ("a1", 1.0), ("ad", 0.1 * 1e-4 * N * N)):
                c = 4.0 if tag == "a1" else 1 + 4 * a
                x = G[t + "lin_solve_b%d_%s_in_x" % (b, tag)].copy()
                oracle.lin_solve(N, b, x, G[t + "lin_solve_b%d_%s_in_x0" % (b, tag)], np.float32(a), np.float32(c), 20)
                assert_bits(x, G[t + "lin_solve_b%d_%s_out" % (b, tag)], "lin_solve")
            for tag, diff in (("d0", 0.0), ("d1", 1e-3)):
                x, x0 = G[t + "blur_b%d_%s_in_x" % (b, tag)].copy(), G[t + "blur_b%d_%s_in_x0" % (b, tag)].copy()
                oracle.diffuse_blur(N, b, x, x0, diff, 0.1, 20)
                assert_bits(x, G[t + "blur_b%d_%s_out_x" % (b, tag)], "blur x")
                assert_bits(x0, G[t + "blur_b%d_%s_out_x0" % (b, tag)], "blur x0")
            for tag in ("slow", "fast"):
                k = t + "advect_b%d_%s_" % (b, tag)
                for mode, name in ((op.ADV_MIDPOINT, "midpoint"), (op.ADV_EULER, "euler")):
                    d = G[k + "in_d"].copy()
                    oracle.advect(N, b, d, G[k + "in_d0"], G[k + "in_u"], G[k + "in_v"], 0.1, mode)
                    assert_bits(d, G[k + "out_" + name], "advect " + name)
        a = [G[t + "project_in_" + k].copy() for k in ("u", "v", "p", "div")]
        oracle.project(N, *a, 20)
        for k, arr in zip(("u", "v", "p", "div"), a):
            assert_bits(arr, G[t + "project_out_" + k], "project " + k)
        f = [G[t + "step_in_%d" % w].copy() for w in range(6)]
        oracle.vel_step(N, f[0], f[1], f[3], f[4], 1e-3, 0.1)
        oracle.dens_step(N, f[2], f[5], f[0], f[1], 1e-3, 0.1)
        for w in range(6):
            assert_bits(f[w], G[t + "step_out_%d" % w], "vel+dens step field %d" % w)


def test_oracle_matches_live_reference(oracle, ref):
    """Fresh random inputs against the unmodified reference, odd sizes included."""
    import oracle_py as op
    rng = np.random.default_rng(7)
    for N in (1, 2, 3, 31, 50):
        for b in (0, 1, 2):
            a = np.float32(0.37)
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            y = x.copy()
            oracle.lin_solve(N, b, x, x0, a, 1 + 4 * a, 20)
            ref.L.ref_lin_solve(N, b, y, x0, a, 1 + 4 * a, 20)
            assert_bits(x, y, "lin_solve N=%d" % N)
            d, d0, u, v = (rand_field(rng, N), rand_field(rng, N), rand_field(rng, N, 3.0), rand_field(rng, N, 3.0))
            e = d.copy()
            oracle.advect(N, b, d, d0, u, v, 0.1, op.ADV_MIDPOINT)
            ref.L.ref_advect(N, b, e, d0, u, v, 0.1)
            assert_bits(d, e, "advect N=%d" % N)
    N = 40
    sim = ref.sim(N, 2e-4, 3e-4, 0.3, 20, 0.05)
    src = op.stam_sources(N) + op.blower_sources(N)
    sim.set_sources(src)
    sim.update(0.05, 7)
    f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.update(N, f, src, 2e-4, 3e-4, 0.05, 7, vorticity=0.3, damp=0.05)
    for name, w in (("u", 0), ("v", 1), ("d", 2), ("u0", 3), ("v0", 4), ("d0", 5)):
        assert_bits(f[name], sim.field(w), "update " + name)
    sim.close()


def test_threaded_oracle_is_bit_identical_to_the_serial_one(oracle):
    """The oracle's order-independent loops (red-black colours, Blur passes, advect, divergence, gradient,
    add_source) may run on several threads so that whole frames at the benchmark size can be checked in
    seconds; every cell is the same expression of the same inputs, so nothing may change."""
    import oracle_py as op
    for N, steps in ((37, 6), (200, 3), (256, 2)):
        src = op.stam_sources(N) + [(0, 3, N - 2, 0, -4.0), (2, N // 3, 5, 1, 0.7)]
        runs = []
        for threads in (1, 7):
            oracle.set_threads(threads)
            try:
                f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
                oracle.update(N, f, src, 1e-4, 1e-4, 0.1, steps, order=op.GS_RB)
            finally:
                oracle.set_threads(1)
            runs.append(f)
        for k in runs[0]:
            assert runs[0][k].tobytes() == runs[1][k].tobytes(), (N, k)


def test_set_bnd_sign_table(oracle):
    """SetBorder sign table and corners (FluidSystem.cpp:465-500)."""
    N = 5
    rng = np.random.default_rng(3)
    for b in (0, 1, 2):
        x = rand_field(rng, N)
        oracle.set_bnd(N, b, x)
        g = x.reshape(N + 2, N + 2)  # g[j, i]
        sx, sy = (-1 if b == 1 else 1), (-1 if b == 2 else 1)
        assert np.array_equal(g[1:-1, 0], sx * g[1:-1, 1]) and np.array_equal(g[1:-1, -1], sx * g[1:-1, -2])
        assert np.array_equal(g[0, 1:-1], sy * g[1, 1:-1]) and np.array_equal(g[-1, 1:-1], sy * g[-2, 1:-1])
        assert g[0, 0] == np.float32(0.5) * (g[0, 1] + g[1, 0])
        assert g[-1, -1] == np.float32(0.5) * (g[-1, -2] + g[-2, -1])


def test_iteration_count_logic(oracle):
    """a == 0 means one sweep (FluidSystem.cpp:538); Blur makes the count odd (:726-727)."""
    N = 8
    rng = np.random.default_rng(5)
    x0 = rand_field(rng, N)
    x1, x20 = rand_field(rng, N), None
    x20 = x1.copy()
    oracle.lin_solve(N, 0, x1, x0, 0.0, 1.0, 1)
    oracle.lin_solve(N, 0, x20, x0, 0.0, 1.0, 20)
    assert_bits(x1, x20, "a=0 ignores iters")
    # blur with iters=20 equals blur with iters=21 (20 is made odd), differs from 19
    a, b_, c_ = rand_field(rng, N), None, None
    s = rand_field(rng, N)
    r20, r21, r19 = a.copy(), a.copy(), a.copy()
    oracle.diffuse_blur(N, 0, r20, s.copy(), 1e-2, 0.1, 20)
    oracle.diffuse_blur(N, 0, r21, s.copy(), 1e-2, 0.1, 21)
    oracle.diffuse_blur(N, 0, r19, s.copy(), 1e-2, 0.1, 19)
    assert_bits(r20, r21, "blur 20 -> 21 passes")
    assert r19.tobytes() != r21.tobytes()
    assert not r20.reshape(N + 2, N + 2)[0].any() and not r20.reshape(N + 2, N + 2)[:, 0].any()  # ZeroBorder


def test_redblack_converges_to_lexicographic(oracle):
    """The fast path's ordering solves the same system: run to convergence the two
    orders agree to <= 1e-4 relative L2 (north_star tolerance; SURVEY 7 hard part 2)."""
    import oracle_py as op
    N = 32
    rng = np.random.default_rng(11)
    x0 = rand_field(rng, N)
    a = np.float32(0.1 * 1e-4 * N * N)
    lex, rb = np.zeros_like(x0), np.zeros_like(x0)
    oracle.lin_solve(N, 1, lex, x0, a, 1 + 4 * a, 400, op.GS_LEX)
    oracle.lin_solve(N, 1, rb, x0, a, 1 + 4 * a, 400, op.GS_RB)
    rel = np.linalg.norm(lex.astype(np.float64) - rb) / np.linalg.norm(lex.astype(np.float64))
    assert rel <= 1e-4, rel


def test_projection_reduces_divergence(oracle):
    N = 32
    # smooth field: the collocated central-difference projection cannot remove grid-scale noise
    jj, ii = np.meshgrid(np.arange(N + 2), np.arange(N + 2), indexing="ij")
    u = (np.sin(2 * np.pi * ii / N) * np.cos(np.pi * jj / N)).astype(np.float32).ravel()
    v = (np.cos(np.pi * ii / N) * np.sin(3 * np.pi * jj / N)).astype(np.float32).ravel()
    oracle.set_bnd(N, 1, u)
    oracle.set_bnd(N, 2, v)

    def div_norm(u, v):
        U, V = u.reshape(N + 2, N + 2).astype(np.float64), v.reshape(N + 2, N + 2).astype(np.float64)
        d = (U[1:-1, 2:] - U[1:-1, :-2]) + (V[2:, 1:-1] - V[:-2, 1:-1])
        return np.linalg.norm(d)

    before = div_norm(u, v)
    p, dv = np.zeros_like(u), np.zeros_like(u)
    oracle.project(N, u, v, p, dv, 200)
    assert div_norm(u, v) < 0.2 * before


def _visc_avg(fluids):
    """visc_avg as the reference computes it (FluidSystem.cpp:26-30): float sum, then float division."""
    v = np.float32(0)
    for _, visc in fluids:
        v = np.float32(v + np.float32(visc))
    return np.float32(v / np.float32(len(fluids)))


def multifluid_expected(oracle, N, steps, fluids, sources, order=0):
    """On the ORIGINAL pipeline the fluids only interact through visc_avg (DensStep1 loops over them,
    FluidSystem.cpp:289-299; VelStep1 never reads a density), so fluid k's density is what a one-fluid
    run with visc = visc_avg, diff = diff_k and fluid k's density sources produces."""
    va = _visc_avg(fluids)
    out = {}
    for k, (diff, _) in enumerate(fluids):
        src = [(2 if s[0] >= 2 else int(s[0]), int(s[1]), int(s[2]), int(s[3]), float(s[4])) for s in sources
               if s[0] < 2 or int(s[0]) == 2 + k]
        f = {n: np.zeros((N + 2) ** 2, np.float32) for n in "u v d u0 v0 d0".split()}
        oracle.update(N, f, src, float(va), float(diff), 0.1, steps, order=order)
        out["d%d" % k] = f["d"]
        out["u"], out["v"] = f["u"], f["v"]
    return out


def test_multifluid_golden_from_the_reference(oracle):
    import os
    from conftest import GOLDEN
    G = np.load(os.path.join(GOLDEN, "golden_multifluid.npz"))
    N, steps = int(G["N"]), int(G["steps"])
    fluids = [tuple(map(float, row)) for row in G["fluids"]]
    want = multifluid_expected(oracle, N, steps, fluids, G["sources"])
    for k in ("u", "v", "d0", "d1"):
        assert_bits(want[k], G[k], "two fluids, " + k)


# ---- FSType::ORIGINAL_BORDERED: the cellInfo-masked pipeline ------------------------------

def _cell_map_with_objects(oracle, N, rng):
    """A cell map with the shipped scene's internal wall plus blocks of non-FLUID (OBJECT) cells, direction
    bits recomputed the way CalcBorderFromContent does after a collider has marked its cells."""
    info = oracle.make_cell_info(N, square_border=False)
    m = info.reshape(N + 2, N + 2)
    for _ in range(3):
        i0, j0 = rng.integers(2, N - 6, 2)
        m[j0:j0 + 4, i0:i0 + 5] = (m[j0:j0 + 4, i0:i0 + 5] & ~np.uint8(16)) | np.uint8(32)
    oracle.calc_border(N, info)
    return info


def test_bordered_cell_map_and_primitives_match_live_reference(oracle, ref):
    rng = np.random.default_rng(77)
    for N in (5, 16, 41):
        for square in (False, True):
            want = np.empty((N + 2) ** 2, np.uint8)
            ref.L.ref_make_cell_info(N, 1 if square else 0, want)
            assert oracle.make_cell_info(N, square).tobytes() == want.tobytes(), (N, square)
        if N > 10:
            info = _cell_map_with_objects(oracle, N, rng)
            chk = info.copy()
            ref.L.ref_calc_border(N, chk)  # idempotent on a map it has produced itself
            assert chk.tobytes() == info.tobytes()
            oracle.L.oracle_square_border(N, info.ctypes.data_as(__import__("ctypes").c_void_p))  # plus the internal wall
        else:
            info = oracle.make_cell_info(N, True)
        r = lambda s=1.0: (rng.uniform(-1, 1, (N + 2) ** 2) * s).astype(np.float32)
        for b in (0, 1, 2):
            for a, c in ((np.float32(0.0), np.float32(1.0)), (np.float32(1.0), np.float32(4.0)), (np.float32(0.37), np.float32(2.48))):
                x, x0 = r(), r()
                y = x.copy()
                oracle.lin_solve_b(N, b, x, x0, a, c, info)
                ref.L.ref_lin_solve_b(N, b, y, x0, a, c, 20, info)
                assert x.tobytes() == y.tobytes(), ("lin_solve_b", N, b, a)
            for diff in (0.0, 2e-3):
                x, x0 = r(), r()
                y, y0 = x.copy(), x0.copy()
                oracle.diffuse_blur_b(N, b, x, x0, diff, 0.1, info)
                ref.L.ref_diffuse_blur_b(N, b, y, y0, diff, 0.1, 20, info)
                assert x.tobytes() == y.tobytes() and x0.tobytes() == y0.tobytes(), ("blur_b", N, b, diff)
        u, v, p, dv = r(), r(), r(), r()
        U, V, P, D = u.copy(), v.copy(), p.copy(), dv.copy()
        oracle.project_b(N, u, v, p, dv, info)
        ref.L.ref_project_b(N, U, V, P, D, 20, info)
        for got, wantf, name in ((u, U, "u"), (v, V, "v"), (p, P, "p"), (dv, D, "div")):
            assert got.tobytes() == wantf.tobytes(), ("project_b", N, name)
        d, d0, uu, vv = r(), r(), r(3.0), r(3.0)
        e = d.copy()
        oracle.advect_b(N, d, d0, uu, vv, 0.1)
        ref.L.ref_advect_b(N, 0, e, d0, uu, vv, 0.1)
        assert d.tobytes() == e.tobytes(), ("advect_b", N)


def test_bordered_frames_match_live_reference(oracle, ref):
    """Whole frames of FluidSystem::Update with FSType::ORIGINAL_BORDERED (VelStep1B + DensStep1BS), with and
    without the scene's SquareBorder, vorticity and damping: the restatement against the unmodified reference,
    including the reference's own quirk that the x velocity receives the y source (FluidSystem.cpp:380)."""
    import oracle_py as op
    for N, steps, square, vort, damp in ((16, 6, True, 0.0, 0.0), (64, 12, True, 0.0, 0.0), (50, 8, False, 0.55, 0.1), (64, 5, True, 0.3, 0.0)):
        c = N // 2
        src = op.stam_sources(N) + [(0, c + 3, c - 2, 0, 4.0), (1, 3, 3, 0, -6.0), (2, c - N // 4, c, 1, 0.8), (0, 5, N - 3, 2, 0.7)]
        sim = ref.sim(N, 1e-4, 2e-4, vort, 20, damp, fstype=1, square_border=square)
        sim.set_sources(src)
        sim.update(0.1, steps)
        info = oracle.make_cell_info(N, square)
        assert info.tobytes() == sim.cell_info().tobytes()
        f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
        oracle.update_b(N, f, info, src, 1e-4, 2e-4, 0.1, steps, vorticity=vort, damp=damp)
        for name, w in (("u", 0), ("v", 1), ("d", 2), ("u0", 3), ("v0", 4), ("d0", 5)):
            assert f[name].tobytes() == sim.field(w).tobytes(), (N, square, name)
        assert np.abs(f["u"]).max() > 0 and np.abs(f["d"]).max() > 0
        sim.close()


def test_multifluid_bordered_primitives_and_frames_match_live_reference(oracle, ref):
    """FSType::ORIGINAL_BORDERED_MF = VelStep1BSO + DensStep1BS, what the shipped scene runs (SystemManager.cpp:103-108):
    LinearSolve1BSO, LinearSolve1BO, ProjectBO and whole frames with two fluids against the unmodified reference."""
    import ctypes
    import oracle_py as op
    rng = np.random.default_rng(4040)
    for N in (7, 24, 45):
        info = _cell_map_with_objects(oracle, N, rng) if N > 10 else oracle.make_cell_info(N, True)
        r = lambda s=1.0: (rng.uniform(-1, 1, (N + 2) ** 2) * s).astype(np.float32)
        for b in (0, 1, 2):
            x, x0 = r(), r()
            y = x.copy()
            oracle.lin_solve_bo(N, b, x, x0, np.float32(0.7), np.float32(3.8), info)
            ref.L.ref_lin_solve_bo(N, b, y, x0, 0.7, 3.8, 20, info)
            assert x.tobytes() == y.tobytes(), ("lin_solve_bo", N, b)
            dens = [np.abs(r(0.6)), np.abs(r(0.9))]
            dens[0].reshape(N + 2, N + 2)[3:5, :] = 0.0
            dens[1].reshape(N + 2, N + 2)[3:6, :] = 0.0   # cells with no fluid at all: the air fallback
            coef = np.array([1e-4, 3e-4], np.float32)
            x, x0 = r(), r()
            y = x.copy()
            aSc = np.float32(0.1) * np.float32(N) * np.float32(N)
            oracle.lin_solve_bso(N, b, x, x0, aSc, np.float32(0.001), dens, coef, info)
            ptrs = (ctypes.c_void_p * 2)(*[d.ctypes.data for d in dens])
            ref.L.ref_lin_solve_bso(N, b, y, x0, aSc, 0.001, 2, ctypes.cast(ptrs, ctypes.c_void_p), coef, 20, info)
            assert x.tobytes() == y.tobytes(), ("lin_solve_bso", N, b)
        u, v, p, dv = r(), r(), r(), r()
        U, V, P, D = u.copy(), v.copy(), p.copy(), dv.copy()
        oracle.project_bo(N, u, v, p, dv, info)
        ref.L.ref_project_bo(N, U, V, P, D, 20, info)
        for got, want, name in ((u, U, "u"), (v, V, "v"), (p, P, "p"), (dv, D, "div")):
            assert got.tobytes() == want.tobytes(), ("project_bo", N, name)
    fluids = [(0.000001, 0.0001), (0.0001, 0.0003)]  # (diff, visc) of the shipped scene
    for N, steps, square, vort, damp in ((16, 5, True, 0.0, 0.0), (48, 10, True, 0.55, 0.1), (64, 6, False, 0.0, 0.0)):
        c = N // 2
        src = [(2, c, c, 0, 100.0), (3, c - 5, c + 4, 0, 60.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0), (3, 4, 4, 1, 0.9), (1, 3, N - 2, 0, -4.0)]
        sim = ref.sim(N, 0, 0, vort, 20, damp, fluids=fluids, fstype=2, square_border=square)
        sim.set_sources(src)
        sim.update(0.1, steps)
        info = oracle.make_cell_info(N, square)
        n = (N + 2) ** 2
        f = {k: np.zeros(n, np.float32) for k in "u v u0 v0 p".split()}
        f["d"] = [np.zeros(n, np.float32) for _ in fluids]
        f["d0"] = [np.zeros(n, np.float32) for _ in fluids]
        oracle.update_mf(N, f, info, src, fluids, 0.1, steps, visc_air=ref.L.ref_visc_air(sim.h), vorticity=vort, damp=damp)
        assert f["u"].tobytes() == sim.field(0).tobytes() and f["v"].tobytes() == sim.field(1).tobytes(), (N, "velocity")
        assert f["u0"].tobytes() == sim.field(3).tobytes() and f["v0"].tobytes() == sim.field(4).tobytes(), (N, "scratch")
        assert f["p"].tobytes() == sim.p().tobytes(), (N, "p")
        for k in range(len(fluids)):
            assert f["d"][k].tobytes() == sim.density(k).tobytes(), (N, "density", k)
        assert np.abs(f["u"]).max() > 0
        sim.close()
