"""Last file pytest collects: the entry point for callers that must not allocate.

`fluid_lin_solve_rb_static` runs the red-black lin_solve on device arrays the
caller already owns and keeps its tile scratch in an array compiled into the
library, so it never calls cudaMalloc/cudaFree.  That is what a real-time loop,
or a plug-in confined to somebody else's memory arena, needs: no allocation
that can fail late or stall the device in the middle of a frame.

The check: its output must equal, bit for bit (the tolerance every red-black
test in this suite uses), what the ordinary entry point `fluid_lin_solve`
produces for the same inputs -- on two calls issued back to back, of different
sizes, so the second one finds the first one's leftovers in the shared array --
and also what the CPU red-black statement (the oracle) produces, so that a
mistake shared by both library paths cannot hide.  A further test watches the
device's free memory across a run of calls: it must not move.
The device memory handed in is allocated by torch here, in the test, never by
the entry point.  Runs alone in a fresh process or after the whole suite.
"""
import ctypes

import numpy as np
import pytest

from conftest import assert_bits, rand_field

pytestmark = pytest.mark.gpu

XO = 31


def pitch_of(N):
    return ((N + 2 + XO + 31) // 32) * 32


def to_device(dense, N):
    import torch
    dev = torch.zeros((N + 2, pitch_of(N)), dtype=torch.float32, device="cuda")
    dev[:, XO:XO + N + 2] = torch.from_numpy(dense.reshape(N + 2, N + 2)).cuda()
    return dev


def to_host(dev, N):
    return dev[:, XO:XO + N + 2].contiguous().cpu().numpy().ravel()


def expected(fl, N, b, x, x0, a, c, iters):
    fl.set_modes(gs_order=fl.GS_REDBLACK, gs_iters=iters)
    y = x.copy()
    fl.lin_solve(N, b, y, x0, a, c)
    fl.set_modes()
    return y


def test_two_calls_in_a_row_match_the_allocating_entry_point(fl, oracle):
    import oracle_py as op
    import torch
    L = fl.lib()
    rng = np.random.default_rng(4242)
    cases = [(1024, 1, np.float32(10.49), 20), (130, 2, np.float32(1.0), 5)]  # (N, b, a, sweeps); c = 1 + 4a
    host, dev = [], []
    for N, b, a, iters in cases:
        x, x0 = rand_field(rng, N), rand_field(rng, N)
        host.append((x, x0))
        dev.append((to_device(x, N), to_device(x0, N)))
    torch.cuda.synchronize()
    for (N, b, a, iters), (dx, dx0) in zip(cases, dev):  # back to back, nothing in between
        rc = L.fluid_lin_solve_rb_static(N, b, dx.data_ptr(), dx0.data_ptr(), a, 1 + 4 * a, iters, None)
        assert rc == 0, L.fluid_last_error()
    torch.cuda.synchronize()
    for (N, b, a, iters), (x, x0), (dx, dx0) in zip(cases, host, dev):
        assert_bits(to_host(dx, N), expected(fl, N, b, x, x0, a, 1 + 4 * a, iters), "static scratch variant, N=%d" % N)
        assert_bits(to_host(dx0, N), x0, "x0 must only be read")
        want = x.copy()
        oracle.lin_solve(N, b, want, x0, a, 1 + 4 * a, iters, op.GS_RB)
        assert_bits(to_host(dx, N), want, "static scratch variant vs the CPU red-black statement, N=%d" % N)


def test_a_run_of_calls_leaves_free_device_memory_unchanged(fl):
    """The point of the entry point: once its first call on the device has done the set-up, no call
    allocates or frees anything.  cudaMemGetInfo before and after a run of calls of different sizes
    and sweep counts (tile passes, single-colour passes) must report the same free byte count."""
    import torch
    L = fl.lib()
    rng = np.random.default_rng(5)
    cases = [(1024, 0, np.float32(1.0), 20), (96, 1, np.float32(0.7), 7), (300, 2, np.float32(0.2), 1), (513, 3, np.float32(0.2), 4)]
    dev = [(to_device(rand_field(rng, N), N), to_device(rand_field(rng, N), N)) for N, _, _, _ in cases]
    rc = L.fluid_lin_solve_rb_static(64, 0, dev[1][0].data_ptr(), dev[1][1].data_ptr(), 1.0, 4.0, 2, None)  # set-up call
    assert rc == 0, L.fluid_last_error()
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        for (N, b, a, iters), (dx, dx0) in zip(cases, dev):
            rc = L.fluid_lin_solve_rb_static(N, b, dx.data_ptr(), dx0.data_ptr(), a, 1 + 4 * a, iters, None)
            assert rc == 0, L.fluid_last_error()
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, (free_before, free_after)


def test_single_sweep_and_odd_counts(fl):
    import torch
    L = fl.lib()
    rng = np.random.default_rng(77)
    for N, b, a, iters in ((64, 0, np.float32(0.0), 20), (200, 1, np.float32(0.3), 1), (257, 2, np.float32(0.3), 7)):
        x, x0 = rand_field(rng, N), rand_field(rng, N)
        dx, dx0 = to_device(x, N), to_device(x0, N)
        rc = L.fluid_lin_solve_rb_static(N, b, dx.data_ptr(), dx0.data_ptr(), a, 1 + 4 * a, iters, None)
        assert rc == 0, L.fluid_last_error()
        torch.cuda.synchronize()
        assert_bits(to_host(dx, N), expected(fl, N, b, x, x0, a, 1 + 4 * a, iters), "N=%d iters=%d" % (N, iters))


def test_sizes_beyond_the_built_in_scratch_are_refused(fl):
    import torch
    L = fl.lib()
    N = 1025
    d = torch.zeros((N + 2, pitch_of(N)), dtype=torch.float32, device="cuda")
    e = torch.zeros_like(d)
    rc = L.fluid_lin_solve_rb_static(N, 0, d.data_ptr(), e.data_ptr(), 1.0, 4.0, 20, None)
    assert rc == 1
    msg = L.fluid_last_error().decode()
    assert "1024" in msg and "N=1025" in msg
    assert L.fluid_lin_solve_rb_static(64, 0, d.data_ptr(), d.data_ptr(), 1.0, 4.0, 20, None) == 1  # aliasing
    assert L.fluid_lin_solve_rb_static(64, 0, d.data_ptr() + 4, e.data_ptr(), 1.0, 4.0, 20, None) == 1  # misaligned
