"""FSType::ORIGINAL_BORDERED on the GPU (first slice): VelStep1B + DensStep1BS and their
cellInfo-masked primitives (FluidSystem.cpp:301-311, 378-398, 550-578, 745-777, 864-874, 1009-1059, 1185-1232,
1388-1408) through the C ABI, bit for bit against

* the oracle restatement (itself pinned to the unmodified reference in tests/test_oracle.py),
* the unmodified reference run live (oracle/_ref, when its prebuilt library travelled with the snapshot),
* golden hashes of reference runs committed under tests/golden/.
"""
import numpy as np
import pytest

from conftest import assert_bits, rand_field

pytestmark = pytest.mark.gpu


def object_map(oracle, N, rng, square):
    """Internal wall + blocks of OBJECT cells, direction bits as CalcBorderFromContent leaves them."""
    info = oracle.make_cell_info(N, False)
    m = info.reshape(N + 2, N + 2)
    for _ in range(3):
        i0, j0 = rng.integers(2, max(3, N - 6), 2)
        m[j0:j0 + 4, i0:i0 + 5] = (m[j0:j0 + 4, i0:i0 + 5] & ~np.uint8(16)) | np.uint8(32)
    m[0, :] &= ~np.uint8(16 | 32)
    oracle.calc_border(N, info)
    if square:
        import ctypes
        oracle.L.oracle_square_border(N, info.ctypes.data_as(ctypes.c_void_p))
    return info


@pytest.mark.parametrize("N", [1, 5, 16, 64, 200, 1030])
def test_cell_map_matches_the_reference_bookkeeping(fl, oracle, N):
    fs = fl.FluidSystem().Setup(N, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    assert fs.GetInfo().tobytes() == oracle.make_cell_info(N, False).tobytes()
    if N >= 5:
        fs.AddSquareBorder()
        assert fs.GetInfo().tobytes() == oracle.make_cell_info(N, True).tobytes()
        fs.ResetCellInfo()
        assert fs.GetInfo().tobytes() == oracle.make_cell_info(N, False).tobytes()
        # a caller-made map (what a collider leaves): FLUID / OBJECT bits uploaded, direction bits recomputed on the device
        rng = np.random.default_rng(N)
        want = object_map(oracle, N, rng, False)
        raw = want.copy()
        raw.reshape(N + 2, N + 2)[1:-1, 1:-1] |= np.uint8(15)  # direction bits deliberately stale
        fs.SetCellInfo(raw, recalc=True)
        assert fs.GetInfo().tobytes() == want.tobytes()
    with pytest.raises(fl.FluidError):
        fs.DisableEdges([(1, 1, 3, 1)])  # not adjacent
    fs.close()
    plain = fl.FluidSystem().Setup(max(N, 2))
    with pytest.raises(fl.FluidError):
        plain.ResetCellInfo()  # an ORIGINAL context has no cell map
    plain.close()


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("N", [3, 16, 65, 200, 300])
def test_bordered_primitives_bitwise(fl, oracle, N, order):
    rng = np.random.default_rng(900 + N)
    info = object_map(oracle, N, rng, N >= 16)
    fs = fl.FluidSystem().Setup(N, gs_order=order, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    fs.SetCellInfo(info)
    # LinearSolve1B
    for b in (0, 1, 2):
        for a, c in ((np.float32(0.0), np.float32(1.0)), (np.float32(1.0), np.float32(4.0)), (np.float32(0.41), np.float32(2.64))):
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            fs.upload(fl.U, x)
            fs.upload(fl.V, x0)
            fs.lin_solve(b, fl.U, fl.V, a, c)
            oracle.lin_solve_b(N, b, x, x0, a, c, info, 20, order)
            assert_bits(fs.download(fl.U), x, "LinearSolve1B N=%d b=%d a=%g order=%d" % (N, b, a, order))
    # BlurB: result and penultimate iterate
    for b in (0, 1):
        for a in (np.float32(0.0), np.float32(0.3)):
            x, x0 = rand_field(rng, N), rand_field(rng, N)
            fs.upload(fl.U, x)
            fs.upload(fl.V, x0)
            fs.blur(b, fl.U, fl.V, a, 1 + 4 * a)
            oracle.blur_b(N, b, x, x0, a, 1 + 4 * a, info)
            assert_bits(fs.download(fl.U), x, "BlurB x N=%d b=%d a=%g" % (N, b, a))
            assert_bits(fs.download(fl.V), x0, "BlurB x0 N=%d b=%d a=%g" % (N, b, a))
    # ProjectB: u, v, p, div
    u, v, p, dv = (rand_field(rng, N) for _ in range(4))
    for fid, arr in ((fl.U, u), (fl.V, v), (fl.U_PREV, p), (fl.V_PREV, dv)):
        fs.upload(fid, arr)
    fs.project(fl.U, fl.V, fl.U_PREV, fl.V_PREV)
    oracle.project_b(N, u, v, p, dv, info, 20, order)
    for fid, arr, name in ((fl.U, u, "u"), (fl.V, v, "v"), (fl.U_PREV, p, "p"), (fl.V_PREV, dv, "div")):
        assert_bits(fs.download(fid), arr, "ProjectB %s N=%d order=%d" % (name, N, order))
    # Advect1B: no SetBorder -- the border cells of the destination stay what they were (both kernels)
    for kernel in (1, 2):
        fs.SetAdvectKernel(kernel)
        d, d0, uu, vv = rand_field(rng, N), rand_field(rng, N), rand_field(rng, N, 0.08), rand_field(rng, N, 0.08)
        for fid, arr in ((fl.DENS, d), (fl.DENS_PREV, d0), (fl.U, uu), (fl.V, vv)):
            fs.upload(fid, arr)
        fs.advect(1, fl.DENS, fl.DENS_PREV, fl.U, fl.V, 0.1)
        oracle.advect_b(N, d, d0, uu, vv, 0.1)
        assert_bits(fs.download(fl.DENS), d, "Advect1B N=%d kernel=%d" % (N, kernel))
        un, vn = rand_field(rng, N), rand_field(rng, N)
        fs.upload(fl.U_PREV, un)
        fs.upload(fl.V_PREV, vn)
        fs.advect_velocity(fl.U_PREV, fl.V_PREV, fl.U, fl.V, 0.1)
        oracle.advect_b(N, un, uu, uu, vv, 0.1)
        oracle.advect_b(N, vn, vv, uu, vv, 0.1)
        assert_bits(fs.download(fl.U_PREV), un, "Advect1B pair u N=%d kernel=%d" % (N, kernel))
        assert_bits(fs.download(fl.V_PREV), vn, "Advect1B pair v N=%d kernel=%d" % (N, kernel))
    fs.close()


def _frame_sources(op, N):
    c = N // 2
    return op.stam_sources(N) + [(0, c + 3, c - 2, 0, 4.0), (1, 3, 3, 0, -6.0), (2, c - N // 4, c, 1, 0.8), (0, 5, N - 3, 2, 0.7)]


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("N,steps,square,vort,damp", [(16, 6, True, 0.0, 0.0), (64, 12, True, 0.0, 0.0), (200, 5, True, 0.0, 0.0),
                                                      (50, 8, False, 0.55, 0.1), (300, 3, True, 0.3, 0.05)])
def test_bordered_frames_vs_oracle_bitwise(fl, oracle, N, steps, square, vort, damp, order):
    """Whole frames (Update with VelStep1B + DensStep1BS), all six arrays, both relaxation orders."""
    import oracle_py as op
    src = _frame_sources(op, N)
    info = oracle.make_cell_info(N, square)
    f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
    oracle.update_b(N, f, info, src, 1e-4, 2e-4, 0.1, steps, order=order, vorticity=vort, damp=damp)
    fs = fl.FluidSystem().Setup(N, 1e-4, 2e-4, vort, damp, gs_order=order, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    if square:
        fs.AddSquareBorder()
    fs.SetSources(src)
    fs.Update(0.1, steps)
    for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS), ("u0", fl.U_PREV), ("v0", fl.V_PREV), ("d0", fl.DENS_PREV)):
        assert_bits(fs.download(fid), f[name], "bordered frame N=%d order=%d: %s" % (N, order, name))
    fs.close()


@pytest.mark.parametrize("N,steps", [(16, 10), (64, 20), (200, 4)])
def test_bordered_frames_vs_live_reference(fl, ref, N, steps):
    """The same against the UNMODIFIED reference executed here (FSType::ORIGINAL_BORDERED + SquareBorder scene)."""
    import oracle_py as op
    src = _frame_sources(op, N)
    sim = ref.sim(N, 1e-4, 2e-4, 0.0, 20, 0.0, fstype=1, square_border=True)
    sim.set_sources(src)
    sim.update(0.1, steps)
    fs = fl.FluidSystem().Setup(N, 1e-4, 2e-4, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    fs.AddSquareBorder()
    fs.SetSources(src)
    fs.Update(0.1, steps)
    assert fs.GetInfo().tobytes() == sim.cell_info().tobytes()
    for name, fid, w in (("u", fl.U, 0), ("v", fl.V, 1), ("d", fl.DENS, 2)):
        assert_bits(fs.download(fid), sim.field(w), "vs live reference N=%d: %s" % (N, name))
    fs.close()
    sim.close()


def test_bordered_golden_frames(fl, golden_bordered):
    """... and against hashes of reference runs committed under tests/golden/ (no reference build needed)."""
    import oracle_py as op
    for fr in golden_bordered:
        N = fr["N"]
        fs = fl.FluidSystem().Setup(N, fr["visc"], fr["diff"], fr["vorticity"], fr["damp"], fs_type=fl.TYPE_ORIGINAL_BORDERED)
        if fr["square_border"]:
            fs.AddSquareBorder()
        fs.SetSources(_frame_sources(op, N))
        fs.Update(fr["dt"], fr["steps"])
        got = (op.fnv1a64(fs.GetDensities()), op.fnv1a64(fs.GetVelX()), op.fnv1a64(fs.GetVelY()))
        assert got == (fr["dens"], fr["u"], fr["v"]), fr
        fs.close()


def test_bordered_ensemble_and_host_calls(fl, oracle):
    """Members of an ensemble share the cell map and stay independent; fluid_ctx_vel_step / dens_step run the B variants."""
    import oracle_py as op
    N, B = 40, 3
    info = oracle.make_cell_info(N, True)
    fs = fl.FluidSystem().Setup(N, 2e-4, 1e-4, gs_order=fl.GS_REDBLACK, batch=B, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    fs.AddSquareBorder()
    want = []
    for b in range(B):
        src = [(2, 5 + 7 * b, 9 + 3 * b, 0, 100.0), (0, 5 + 7 * b, 9 + 3 * b, 0, 5.0), (1, 20, 20 + b, 0, -5.0)]
        fs.SetSources(src, sim=b)
        f = {k: np.zeros((N + 2) ** 2, np.float32) for k in "u v d u0 v0 d0".split()}
        oracle.update_b(N, f, info, src, 2e-4, 1e-4, 0.1, 4, order=op.GS_RB)
        want.append(f)
    fs.Update(0.1, 4)
    for b in range(B):
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            assert_bits(fs.download(fid, b), want[b][name], "member %d %s" % (b, name))
    fs.close()
    # vel_step / dens_step alone (no zero fill): VelStep1B's first AddArrDt takes the Y source for vX too
    rng = np.random.default_rng(3)
    fs = fl.FluidSystem().Setup(N, 2e-4, 1e-4, fs_type=fl.TYPE_ORIGINAL_BORDERED)
    fs.AddSquareBorder()
    arrs = {k: rand_field(rng, N, 0.2) for k in "u v d u0 v0 d0".split()}
    ids = {"u": fl.U, "v": fl.V, "d": fl.DENS, "u0": fl.U_PREV, "v0": fl.V_PREV, "d0": fl.DENS_PREV}
    for k, a in arrs.items():
        fs.upload(ids[k], a)
    fs.VelStep(0.1)
    fs.DensStep(0.1)
    oracle.vel_step_b(N, arrs["u"], arrs["v"], arrs["u0"], arrs["v0"], 2e-4, 0.1, info)
    oracle.dens_step_bs(N, arrs["d"], arrs["d0"], arrs["u"], arrs["v"], 1e-4, 0.1, info)
    for k, a in arrs.items():
        assert_bits(fs.download(ids[k]), a, "VelStep1B/DensStep1BS: %s" % k)
    fs.close()
    with pytest.raises(fl.FluidError):
        fl.FluidSystem().Setup(128, gs_order=fl.GS_REDBLACK, slab=(0, 2, 64), fs_type=fl.TYPE_ORIGINAL_BORDERED)


# ---- FSType::ORIGINAL_BORDERED_MF: VelStep1BSO + DensStep1BS, what the shipped scene runs -----------------------------------

SCENE_FLUIDS = [(0.000001, 0.0001), (0.0001, 0.0003)]  # (diff, visc), SystemManager.cpp:103-106


def _mf_sources(N):
    c = N // 2
    return [(2, c, c, 0, 100.0), (3, c - 5, c + 4, 0, 60.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0), (3, 4, 4, 1, 0.9), (1, 3, N - 2, 0, -4.0)]


def _mf_context(fl, N, order, square, vort=0.0, damp=0.0, info=None):
    fs = fl.FluidSystem().Setup(N, 0.0, 0.0, vort, damp, gs_order=order, n_fluids=len(SCENE_FLUIDS), fs_type=fl.TYPE_ORIGINAL_BORDERED_MF)
    for k, (diff, visc) in enumerate(SCENE_FLUIDS):
        fs.SetFluid(k, diff, visc)
    if info is not None:
        fs.SetCellInfo(info)
    elif square:
        fs.AddSquareBorder()
    return fs


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("N,steps,square,vort,damp", [(16, 5, True, 0.0, 0.0), (48, 10, True, 0.55, 0.1), (130, 4, False, 0.0, 0.0), (200, 3, True, 0.3, 0.0)])
def test_multifluid_bordered_frames_vs_oracle_bitwise(fl, oracle, N, steps, square, vort, damp, order):
    src = _mf_sources(N)
    info = oracle.make_cell_info(N, square)
    n = (N + 2) ** 2
    f = {k: np.zeros(n, np.float32) for k in "u v u0 v0 p".split()}
    f["d"] = [np.zeros(n, np.float32) for _ in SCENE_FLUIDS]
    f["d0"] = [np.zeros(n, np.float32) for _ in SCENE_FLUIDS]
    oracle.update_mf(N, f, info, src, SCENE_FLUIDS, 0.1, steps, order=order, vorticity=vort, damp=damp)
    fs = _mf_context(fl, N, order, square, vort, damp)
    fs.SetSources([(fl.dens_of(s[0] - 2) if s[0] >= 2 else s[0],) + tuple(s[1:]) for s in src])
    fs.Update(0.1, steps)
    assert_bits(fs.download(fl.U), f["u"], "MF frame N=%d order=%d: u" % (N, order))
    assert_bits(fs.download(fl.V), f["v"], "MF frame N=%d order=%d: v" % (N, order))
    assert_bits(fs.download(fl.U_PREV), f["u0"], "MF frame N=%d order=%d: u.Prev (untouched by ProjectBO)" % (N, order))
    assert_bits(fs.download(fl.V_PREV), f["v0"], "MF frame N=%d order=%d: v.Prev (div)" % (N, order))
    for k in range(len(SCENE_FLUIDS)):
        assert_bits(fs.download(fl.dens_of(k)), f["d"][k], "MF frame N=%d order=%d: density %d" % (N, order, k))
    fs.close()


def test_multifluid_bordered_with_object_cells(fl, oracle):
    """Non-FLUID cells inside the grid carrying a velocity (what a collider leaves): their velocity enters ProjectBO's
    divergence and LinearSolve1BO."""
    N, steps = 60, 3
    rng = np.random.default_rng(8)
    info = object_map(oracle, N, rng, True)
    src = _mf_sources(N)
    n = (N + 2) ** 2
    init_u, init_v = rand_field(rng, N, 0.3), rand_field(rng, N, 0.3)
    for order in (0, 1):
        f = {"u": init_u.copy(), "v": init_v.copy(), "u0": np.zeros(n, np.float32), "v0": np.zeros(n, np.float32), "p": np.zeros(n, np.float32)}
        f["d"] = [np.zeros(n, np.float32) for _ in SCENE_FLUIDS]
        f["d0"] = [np.zeros(n, np.float32) for _ in SCENE_FLUIDS]
        oracle.update_mf(N, f, info, src, SCENE_FLUIDS, 0.1, steps, order=order)
        fs = _mf_context(fl, N, order, False, info=info)
        fs.upload(fl.U, init_u)
        fs.upload(fl.V, init_v)
        fs.SetSources([(fl.dens_of(s[0] - 2) if s[0] >= 2 else s[0],) + tuple(s[1:]) for s in src])
        fs.Update(0.1, steps)
        assert_bits(fs.download(fl.U), f["u"], "objects, order %d: u" % order)
        assert_bits(fs.download(fl.V), f["v"], "objects, order %d: v" % order)
        assert_bits(fs.download(fl.dens_of(1)), f["d"][1], "objects, order %d: density 1" % order)
        fs.close()


@pytest.mark.parametrize("N,steps", [(16, 8), (64, 12), (192, 3)])
def test_multifluid_bordered_frames_vs_live_reference(fl, ref, N, steps):
    """The shipped scene's solver set-up (two fluids, ORIGINAL_BORDERED_MF, SquareBorder, vorticity 0.55, damp 0.1;
    SystemManager.cpp:99-108) against the UNMODIFIED reference executed here."""
    src = _mf_sources(N)
    sim = ref.sim(N, 0, 0, 0.55, 20, 0.1, fluids=SCENE_FLUIDS, fstype=2, square_border=True)
    sim.set_sources(src)
    sim.update(0.1, steps)
    fs = _mf_context(fl, N, 0, True, 0.55, 0.1)
    fs.SetSources([(fl.dens_of(s[0] - 2) if s[0] >= 2 else s[0],) + tuple(s[1:]) for s in src])
    fs.Update(0.1, steps)
    assert_bits(fs.download(fl.U), sim.field(0), "vs live reference N=%d: u" % N)
    assert_bits(fs.download(fl.V), sim.field(1), "vs live reference N=%d: v" % N)
    for k in range(len(SCENE_FLUIDS)):
        assert_bits(fs.download(fl.dens_of(k)), sim.density(k), "vs live reference N=%d: density %d" % (N, k))
    fs.close()
    sim.close()


def test_multifluid_bordered_golden_frames(fl, golden_mf):
    import oracle_py as op
    for fr in golden_mf:
        N = fr["N"]
        fs = _mf_context(fl, N, 0, fr["square_border"], fr["vorticity"], fr["damp"])
        fs.SetSources([(fl.dens_of(s[0] - 2) if s[0] >= 2 else s[0],) + tuple(s[1:]) for s in _mf_sources(N)])
        fs.Update(fr["dt"], fr["steps"])
        got = (op.fnv1a64(fs.download(fl.dens_of(0))), op.fnv1a64(fs.download(fl.dens_of(1))), op.fnv1a64(fs.GetVelX()), op.fnv1a64(fs.GetVelY()))
        assert got == (fr["dens0"], fr["dens1"], fr["u"], fr["v"]), fr
        fs.close()
