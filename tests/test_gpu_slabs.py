"""GPU suite: the multi-GPU row-slab path, exercised on ONE GPU by
running every slab as its own context + host thread (fluid_b200.slabs.LocalSlabGroup;
the exchange points, halo regions and kernels are exactly those a multi-process
NCCL run uses).  Red-black results do not depend on the decomposition, so the
assembled global fields must equal the CPU red-black statement bit for bit."""
import numpy as np
import pytest

from conftest import assert_bits, rand_field

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("count,rows", [(2, 96), (3, 64), (4, 50)])
def test_slabs_equal_single_domain_bitwise(fl, oracle, count, rows):
    import oracle_py as op
    from fluid_b200.slabs import LocalSlabGroup
    N = count * rows  # square global grid so that the oracle can state it
    steps = 3
    rng = np.random.default_rng(count)
    src = op.stam_sources(N) + [(2, 7, N - 3, 0, 50.0), (0, N - 5, 3, 0, -4.0), (1, 9, rows, 0, 3.0), (1, 10, rows + 1, 0, 3.0)]
    init = {k: rand_field(rng, N, 0.5) for k in "u v d".split()}
    f = {k: init[k].copy() for k in init}
    f.update({k: np.zeros((N + 2) ** 2, np.float32) for k in "u0 v0 d0".split()})
    oracle.update(N, f, src, 1e-4, 1e-4, 0.1, steps, order=op.GS_RB)

    grp = LocalSlabGroup(N, count, rows, visc=1e-4, diff=1e-4, gs_order=fl.GS_REDBLACK)
    try:
        grp.SetSources(src)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            grp.upload(fid, init[name])
        grp.Update(0.1, steps)
        assert grp.exchanges == 6 * steps  # 4 exchange points in vel_step (u,v), 2 in dens_step = 10 field transfers
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            out = np.full((N + 2) ** 2, np.nan, np.float32)
            grp.download(fid, out)
            assert_bits(out, f[name], "%d slabs of %d rows: %s" % (count, rows, name))
    finally:
        grp.close()


def test_large_slabs_use_the_streaming_kernels_bitwise(fl, oracle):
    """Two slabs of 2048 x 1024 cells are large enough for the register-resident streaming kernels
    (red-black relaxation and Blur); there the slab's inner end is not a wall -- the row beyond belongs
    to the neighbour and stays as exchanged -- while its outer end is."""
    import oracle_py as op
    from fluid_b200.slabs import LocalSlabGroup
    N, count, rows = 2048, 2, 1024
    rng = np.random.default_rng(77)
    src = op.stam_sources(N) + [(2, 7, N - 3, 0, 50.0), (1, 9, rows, 0, 3.0), (1, 10, rows + 1, 0, 3.0)]
    init = {k: rand_field(rng, N, 0.5) for k in "u v d".split()}
    f = {k: init[k].copy() for k in init}
    f.update({k: np.zeros((N + 2) ** 2, np.float32) for k in "u0 v0 d0".split()})
    oracle.update(N, f, src, 1e-4, 1e-4, 0.1, 1, order=op.GS_RB)
    grp = LocalSlabGroup(N, count, rows, visc=1e-4, diff=1e-4, gs_order=fl.GS_REDBLACK)
    try:
        grp.SetSources(src)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            grp.upload(fid, init[name])
        grp.Update(0.1, 1)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            out = np.full((N + 2) ** 2, np.nan, np.float32)
            grp.download(fid, out)
            assert_bits(out, f[name], "2 slabs of 1024 rows, N=2048: %s" % name)
    finally:
        grp.close()


def test_backtrace_longer_than_the_halo_is_reported(fl):
    from fluid_b200.slabs import LocalSlabGroup
    N, count, rows = 128, 2, 64
    grp = LocalSlabGroup(N, count, rows, visc=0.0, diff=0.0, gs_order=fl.GS_REDBLACK)
    try:
        v = np.full((N + 2) ** 2, 60.0, np.float32)  # dt*N*v = 768 cells per step
        grp.upload(fl.V, v)
        with pytest.raises(fl.FluidError) as e:
            grp.Update(0.1, 1)
        assert "halo" in str(e.value)
    finally:
        grp.close()


def test_slab_argument_checks(fl):
    with pytest.raises(fl.FluidError):
        fl.FluidSystem().Setup(128, gs_order=fl.GS_WAVEFRONT, slab=(0, 2, 64))  # reference order is a pipeline across slabs
    with pytest.raises(fl.FluidError):
        fl.FluidSystem().Setup(128, gs_order=fl.GS_REDBLACK, slab=(0, 2, 40))   # fewer rows than the halo
    with pytest.raises(fl.FluidError):
        fl.FluidSystem().Setup(128, gs_order=fl.GS_REDBLACK, slab=(2, 2, 64))   # index out of range
    fs = fl.FluidSystem().Setup(128, gs_order=fl.GS_REDBLACK, slab=(1, 2, 64))
    with pytest.raises(fl.FluidError) as e:
        fs.Update(0.1, 1)  # no transport registered
    assert "exchange" in str(e.value)
    sd, rd, su, ru, n = fs.HaloRegions(fl.U)
    assert su is None and ru is None and sd and rd and n == 49 * (((128 + 2 + 31 + 31) // 32) * 32)
    # the halo bounds the sweep count (2 rows per red-black sweep + 2), at creation and afterwards alike
    with pytest.raises(fl.FluidError) as e:
        fs.SetModes(fl.GS_REDBLACK, 30)
    assert "halo" in str(e.value)
    with pytest.raises(fl.FluidError):
        fs.SetModes(fl.GS_WAVEFRONT, 20)
    fs.SetModes(fl.GS_REDBLACK, 23)  # 2*23 + 2 = 48 rows: just fits
    fs.close()
    with pytest.raises(fl.FluidError):
        fl.FluidSystem().Setup(128, gs_order=fl.GS_REDBLACK, gs_iters=30, slab=(0, 2, 64))  # 62 rows needed, 48 there
    fl.FluidSystem().Setup(128, gs_order=fl.GS_REDBLACK, gs_iters=30, slab=(0, 2, 64), slab_halo=62).close()


@pytest.mark.parametrize("order,N,rows,count", [(0, 150, 60, 3), (1, 150, 60, 3), (1, 2048, 512, 2)])
def test_non_square_single_context(fl, order, N, rows, count):
    """N x NY grids (what a multi-GPU weak-scaling run is globally).  The oracle is square-only, so
    the check is internal consistency: red-black on one context == the slab decomposition (which is
    itself pinned to the oracle on square grids); the reference order runs and stays finite.
    2048 x 1024: the single context is large enough for the streaming kernels, its two slabs are not,
    so this also sets the two red-black implementations against each other on a non-square grid."""
    from fluid_b200.slabs import LocalSlabGroup
    steps = 2
    NY = rows * count
    rng = np.random.default_rng(31)
    init = {k: (rng.uniform(-1, 1, (N + 2) * (NY + 2)) * 0.4).astype(np.float32) for k in "uvd"}
    src = [(2, N // 2, NY // 2, 0, 100.0), (0, N // 2, NY // 2, 0, 5.0), (1, 7, rows, 0, 5.0)]
    fs = fl.FluidSystem().Setup(N, 1e-4, 1e-4, gs_order=order, slab=(0, 1, NY))
    fs.SetSources(src)
    for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
        fs.upload(fid, init[name])
    fs.Update(0.1, steps)
    one = {name: fs.download(fid) for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS))}
    fs.close()
    assert all(np.isfinite(a).all() and a.size == (N + 2) * (NY + 2) for a in one.values())
    if order == 1:
        grp = LocalSlabGroup(N, count, rows, visc=1e-4, diff=1e-4, gs_order=fl.GS_REDBLACK)
        try:
            grp.SetSources(src)
            for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
                grp.upload(fid, init[name])
            grp.Update(0.1, steps)
            for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
                out = np.zeros((N + 2) * (NY + 2), np.float32)
                grp.download(fid, out)
                assert_bits(out, one[name], "non-square %s: slabs vs one context" % name)
        finally:
            grp.close()
