"""GPU suite: fluid_multi_* -- several GPUs driven from ONE process through the C ABI, halos by peer
copies ordered with CUDA events.  A device may be listed several times, so the whole mechanism
(threads, barriers, events, peer copies) also runs on a single-GPU box; with >= 2 GPUs the same test
spreads the slabs over real devices.  Red-black results are decomposition independent: bitwise check."""
import numpy as np
import pytest

from conftest import assert_bits, rand_field

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("ndev", [2, 3])
def test_multi_device_peer_copy_halos_bitwise(fl, oracle, ndev):
    import torch
    import oracle_py as op
    have = torch.cuda.device_count()
    devices = [k % have for k in range(ndev)]
    rows, steps = 70, 3
    N = rows * ndev
    rng = np.random.default_rng(ndev)
    init = {k: rand_field(rng, N, 0.5) for k in "u v d".split()}
    src = op.stam_sources(N) + [(1, 9, rows, 0, 3.0), (0, 11, rows + 1, 0, -2.0), (2, N - 2, 2 * rows - 1, 1, 0.5)]
    f = {k: init[k].copy() for k in init}
    f.update({k: np.zeros((N + 2) ** 2, np.float32) for k in "u0 v0 d0".split()})
    oracle.update(N, f, src, 1e-4, 2e-4, 0.1, steps, order=op.GS_RB)

    m = fl.FluidMulti(N, rows, devices, visc=1e-4, diff=2e-4)
    try:
        m.SetSources(src)
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            m.upload(fid, init[name])
        m.Update(0.1, steps)
        m.Sync()
        assert m.LastStepMs() > 0
        for name, fid in (("u", fl.U), ("v", fl.V), ("d", fl.DENS)):
            assert_bits(m.download(fid), f[name], "%d slabs on devices %r: %s" % (ndev, devices, name))
    finally:
        m.close()


def test_multi_reports_a_failing_slab(fl):
    m = fl.FluidMulti(128, 64, [0, 0], visc=0.0, diff=0.0)
    try:
        m.upload(fl.V, np.full((128 + 2) ** 2, 60.0, np.float32))  # backtrace far longer than the halo
        m.Update(0.1, 1)
        with pytest.raises(fl.FluidError) as e:
            m.Sync()
        assert "halo" in str(e.value)
    finally:
        m.close()
    with pytest.raises(fl.FluidError):
        fl.FluidMulti(128, 30, [0, 0])  # fewer rows than the halo
