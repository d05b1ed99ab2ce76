"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the two CPU checkers.

* ``Oracle``  -> oracle/liboracle.so, the plain-C restatement (oracle/stam_oracle.c)
* ``Ref``     -> oracle/_ref/libfluid_ref.so, the UNMODIFIED reference solver
                 compiled from /root/reference (oracle/ref_build/ref_capi.cpp)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libfluid_ref.so")

GS_LEX, GS_RB = 0, 1
ADV_MIDPOINT, ADV_EULER = 0, 1

c_int, c_float, c_void_p = ctypes.c_int, ctypes.c_float, ctypes.c_void_p
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile liboracle.so (always possible) and, when /root/reference is
    mounted, oracle/_ref/libfluid_ref.so.  Building the checker is not using it."""
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
            os.path.join(HERE, "stam_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, os.path.join(HERE, "liboracle.so")])
    ref_src = os.environ.get("FLUID_REF", "/root/reference")
    if os.path.exists(os.path.join(ref_src, "FluidSim", "src", "FluidSystem.cpp")):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref", "FLUID_REF=" + ref_src])


def fnv1a64(arr):
    """FNV-1a 64 over the raw bytes of an array (vectorised per byte lane would
    change the value, so this is the plain serial definition on a bytes object)."""
    h = 0xCBF29CE484222325
    for b in np.ascontiguousarray(arr).tobytes():
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


class SourceEntry(ctypes.Structure):
    _fields_ = [("field", c_int), ("i", c_int), ("j", c_int), ("mode", c_int), ("value", c_float)]


def stam_sources(N):
    """Source S1 of SURVEY section 8(d): dens_prev=100, u_prev=v_prev=5 at (N/2,N/2)."""
    c = N // 2
    return [(2, c, c, 0, 100.0), (0, c, c, 0, 5.0), (1, c, c, 0, 5.0)]


def blower_sources(N):
    """Source S2: the reference's Blower(0.01,0.01,1,1,0.01,1,1.8,0) as scene code
    creates it (SystemManager.cpp:91; FluidUpdaters.cpp:45-67), float arithmetic.
    Note px=py=0.01 in the shipped scene (SURVEY 8(d) quotes p=0.5 for a centred one)."""
    f = np.float32
    px = py = f(0.01)
    scale = f(0.01)
    vel = f(1.0) * f(N + 2) / f(64.0)
    lo = int(f(1) + -scale * f(N) + px * f(N))
    hi = int(f(1) + scale * f(N) + px * f(N))
    out = []
    for i in range(lo, hi):
        for j in range(lo, hi):
            out += [(2, i, j, 1, 1.8), (0, i, j, 2, float(vel)), (1, i, j, 2, float(vel))]
    return out


class Oracle:
    def __init__(self, path=ORACLE_SO):
        L = self.L = ctypes.CDLL(path)
        L.oracle_add_source.argtypes = [c_int, _f32p, _f32p, c_float]
        L.oracle_set_bnd.argtypes = [c_int, c_int, _f32p]
        L.oracle_zero_bnd.argtypes = [c_int, _f32p]
        L.oracle_lin_solve.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.oracle_lin_solve_rb.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.oracle_diffuse.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, c_int]
        L.oracle_diffuse_blur.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.oracle_blur.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.oracle_advect.argtypes = [c_int, c_int, _f32p, _f32p, _f32p, _f32p, c_float, c_int]
        L.oracle_project.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_int, c_int]
        L.oracle_vel_step.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_float, c_float, c_int, c_int, c_int]
        L.oracle_dens_step.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_float, c_float, c_int, c_int]
        L.oracle_vort_confinement.argtypes = [c_int, _f32p, _f32p, c_float, _f32p, _f32p, _f32p, _f32p]
        L.oracle_damp.argtypes = [c_int, _f32p, c_float, c_float]
        L.oracle_update.argtypes = [c_int] + [_f32p] * 6 + [c_int, c_void_p, c_float, c_float, c_float, c_int, c_int,
                                                           c_int, c_float, c_float, c_int]
        L.oracle_render.argtypes = [c_int, c_int, c_void_p, c_void_p, _f32p, _f32p, _f32p, _f32p, _f32p]
        L.oracle_total_inner_mass.argtypes = [c_int, _f32p]
        L.oracle_total_inner_mass.restype = c_float
        L.oracle_set_threads.argtypes = [c_int]
        _u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
        L.oracle_make_cell_info.argtypes = [c_int, c_int, _u8p]
        L.oracle_calc_border.argtypes = [c_int, _u8p]
        L.oracle_lin_solve_b.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.oracle_lin_solve_b_rb.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.oracle_blur_b.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.oracle_diffuse_blur_b.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.oracle_project_b.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_int, c_int, _u8p]
        L.oracle_advect_b.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_float, c_int]
        L.oracle_vel_step_b.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_float, c_float, c_int, c_int, c_int, _u8p]
        L.oracle_dens_step_bs.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_float, c_float, c_int, c_int, _u8p]
        L.oracle_lin_solve_bso.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, c_void_p, _f32p, c_int, c_int, _u8p]
        L.oracle_lin_solve_bo.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, c_int, _u8p]
        L.oracle_project_bo.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_int, c_int, _u8p]
        L.oracle_update_mf.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, _f32p, c_int, c_void_p, c_void_p, _f32p, _f32p, c_float, _u8p,
                                       c_int, c_void_p, c_float, c_int, c_int, c_int, c_float, c_float, c_int]
        L.oracle_update_b.argtypes = [c_int] + [_f32p] * 6 + [_u8p, c_int, c_void_p, c_float, c_float, c_float, c_int, c_int,
                                                              c_int, c_float, c_float, c_int]

    def set_threads(self, n):
        """Threads for the order-independent loops (bit-identical results; 1 = the serial restatement)."""
        self.L.oracle_set_threads(n)

    def add_source(self, N, x, s, dt):
        self.L.oracle_add_source(N, x, s, dt)

    def set_bnd(self, N, b, x):
        self.L.oracle_set_bnd(N, b, x)

    def lin_solve(self, N, b, x, x0, a, c, iters=20, order=GS_LEX):
        (self.L.oracle_lin_solve_rb if order == GS_RB else self.L.oracle_lin_solve)(N, b, x, x0, a, c, iters)

    def diffuse(self, N, b, x, x0, diff, dt, iters=20, order=GS_LEX):
        self.L.oracle_diffuse(N, b, x, x0, diff, dt, iters, order)

    def diffuse_blur(self, N, b, x, x0, diff, dt, iters=20):
        self.L.oracle_diffuse_blur(N, b, x, x0, diff, dt, iters)

    def blur(self, N, b, x, x0, a, c, iters=20):
        self.L.oracle_blur(N, b, x, x0, a, c, iters)

    def advect(self, N, b, d, d0, u, v, dt, mode=ADV_MIDPOINT):
        self.L.oracle_advect(N, b, d, d0, u, v, dt, mode)

    def project(self, N, u, v, p, div, iters=20, order=GS_LEX):
        self.L.oracle_project(N, u, v, p, div, iters, order)

    def vel_step(self, N, u, v, u0, v0, visc, dt, iters=20, order=GS_LEX, mode=ADV_MIDPOINT):
        self.L.oracle_vel_step(N, u, v, u0, v0, visc, dt, iters, order, mode)

    def dens_step(self, N, x, x0, u, v, diff, dt, iters=20, mode=ADV_MIDPOINT):
        self.L.oracle_dens_step(N, x, x0, u, v, diff, dt, iters, mode)

    def vort_confinement(self, N, u, v, vorticity):
        n = (N + 2) ** 2
        scr = [np.zeros(n, np.float32) for _ in range(4)]
        self.L.oracle_vort_confinement(N, u, v, vorticity, *scr)

    def damp(self, N, vel, amount, dt):
        self.L.oracle_damp(N, vel, amount, dt)

    def update(self, N, fields, sources, visc, diff, dt, nsteps=1, iters=20, order=GS_LEX, mode=ADV_MIDPOINT,
               vorticity=0.0, damp=0.0):
        """fields = dict(u,v,d,u0,v0,d0) of flat float32 arrays, advanced in place."""
        arr = (SourceEntry * max(1, len(sources)))(*[SourceEntry(*s) for s in sources])
        self.L.oracle_update(N, fields["u"], fields["v"], fields["d"], fields["u0"], fields["v0"], fields["d0"],
                             len(sources), ctypes.cast(arr, c_void_p), visc, diff, dt, iters, order, mode, vorticity,
                             damp, nsteps)

    def total_inner_mass(self, N, d):
        return float(self.L.oracle_total_inner_mass(N, d))

    # ---- FSType::ORIGINAL_BORDERED (cellInfo-masked variants) ----
    def make_cell_info(self, N, square_border=False):
        """cellInfo as Update rebuilds it each frame: ResetCellInfo, CalcBorderFromContent[, SquareBorder]."""
        info = np.empty((N + 2) ** 2, np.uint8)
        self.L.oracle_make_cell_info(N, 1 if square_border else 0, info)
        return info

    def calc_border(self, N, info):
        self.L.oracle_calc_border(N, info)

    def lin_solve_b(self, N, b, x, x0, a, c, info, iters=20, order=GS_LEX):
        (self.L.oracle_lin_solve_b_rb if order == GS_RB else self.L.oracle_lin_solve_b)(N, b, x, x0, a, c, iters, info)

    def blur_b(self, N, b, x, x0, a, c, info, iters=20):
        self.L.oracle_blur_b(N, b, x, x0, a, c, iters, info)

    def diffuse_blur_b(self, N, b, x, x0, diff, dt, info, iters=20):
        self.L.oracle_diffuse_blur_b(N, b, x, x0, diff, dt, iters, info)

    def project_b(self, N, u, v, p, div, info, iters=20, order=GS_LEX):
        self.L.oracle_project_b(N, u, v, p, div, iters, order, info)

    def advect_b(self, N, d, d0, u, v, dt, mode=ADV_MIDPOINT):
        self.L.oracle_advect_b(N, d, d0, u, v, dt, mode)

    def vel_step_b(self, N, u, v, u0, v0, visc, dt, info, iters=20, order=GS_LEX, mode=ADV_MIDPOINT):
        self.L.oracle_vel_step_b(N, u, v, u0, v0, visc, dt, iters, order, mode, info)

    def dens_step_bs(self, N, x, x0, u, v, diff, dt, info, iters=20, mode=ADV_MIDPOINT):
        self.L.oracle_dens_step_bs(N, x, x0, u, v, diff, dt, iters, mode, info)

    # ---- FSType::ORIGINAL_BORDERED_MF ----
    @staticmethod
    def _ptrs(arrays):
        return ctypes.cast((c_void_p * len(arrays))(*[a.ctypes.data for a in arrays]), c_void_p)

    def lin_solve_bso(self, N, b, x, x0, aSc, a_fallback, dens, coef, info, iters=20, order=GS_LEX):
        self.L.oracle_lin_solve_bso(N, b, x, x0, aSc, a_fallback, len(dens), self._ptrs(dens), np.asarray(coef, np.float32), iters, order, info)

    def lin_solve_bo(self, N, b, x, x0, a, c, info, iters=20, order=GS_LEX):
        self.L.oracle_lin_solve_bo(N, b, x, x0, a, c, iters, order, info)

    def project_bo(self, N, u, v, p, div, info, iters=20, order=GS_LEX):
        self.L.oracle_project_bo(N, u, v, p, div, iters, order, info)

    def update_mf(self, N, f, info, sources, fluids, dt, nsteps=1, visc_air=0.001, iters=20, order=GS_LEX, mode=ADV_MIDPOINT,
                  vorticity=0.0, damp=0.0):
        """f: dict with u, v, u0, v0, p and lists d, d0 (one array per fluid); fluids: [(diff, visc), ...]."""
        arr = (SourceEntry * max(1, len(sources)))(*[SourceEntry(*s) for s in sources])
        diff = np.array([x[0] for x in fluids], np.float32)
        visc = np.array([x[1] for x in fluids], np.float32)
        self.L.oracle_update_mf(N, f["u"], f["v"], f["u0"], f["v0"], f["p"], len(fluids), self._ptrs(f["d"]), self._ptrs(f["d0"]), diff, visc,
                                visc_air, info, len(sources), ctypes.cast(arr, c_void_p), dt, iters, order, mode, vorticity, damp, nsteps)

    def update_b(self, N, fields, info, sources, visc, diff, dt, nsteps=1, iters=20, order=GS_LEX, mode=ADV_MIDPOINT,
                 vorticity=0.0, damp=0.0):
        arr = (SourceEntry * max(1, len(sources)))(*[SourceEntry(*s) for s in sources])
        self.L.oracle_update_b(N, fields["u"], fields["v"], fields["d"], fields["u0"], fields["v0"], fields["d0"], info,
                               len(sources), ctypes.cast(arr, c_void_p), visc, diff, dt, iters, order, mode, vorticity,
                               damp, nsteps)

    def render(self, N, dens, u, v, color, info=None):
        """dens: one array or a list of arrays (one per fluid); color: 3 floats per fluid; info: cell map or None."""
        n = (N + 2) ** 2
        if isinstance(dens, np.ndarray):
            dens = [dens]
        dens = [np.ascontiguousarray(d, np.float32) for d in dens]
        ptrs = (c_void_p * len(dens))(*[d.ctypes.data for d in dens])
        px, ln = np.empty(n * 3, np.float32), np.empty(n * 4, np.float32)
        inf = None if info is None else np.ascontiguousarray(info, np.uint8)
        self.L.oracle_render(N, len(dens), ctypes.cast(ptrs, c_void_p), None if inf is None else inf.ctypes.data, u, v,
                             np.ascontiguousarray(np.asarray(color, np.float32).ravel()), px, ln)
        return px.reshape(n, 3), ln.reshape(n, 4)


class Ref:
    """The unmodified reference (FluidSystem, FSType::ORIGINAL, one fluid)."""

    def __init__(self, path=REF_SO):
        L = self.L = ctypes.CDLL(path)
        L.ref_create.restype = c_void_p
        L.ref_create.argtypes = [c_int, c_float, c_float, c_float, c_int, c_int, c_float]
        L.ref_create_multi.restype = c_void_p
        L.ref_create_multi.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_int, c_int, c_float]
        L.ref_density.restype = ctypes.POINTER(c_float)
        L.ref_density.argtypes = [c_void_p, c_int, c_int]
        L.ref_destroy.argtypes = [c_void_p]
        L.ref_set_sources.argtypes = [c_void_p, c_int] + [c_void_p] * 5
        L.ref_field.restype = ctypes.POINTER(c_float)
        L.ref_field.argtypes = [c_void_p, c_int]
        L.ref_update.argtypes = [c_void_p, c_float, c_int]
        L.ref_vel_step.argtypes = [c_void_p, c_float]
        L.ref_dens_step.argtypes = [c_void_p, c_float]
        L.ref_vort_confinement.argtypes = [c_void_p]
        L.ref_damp.argtypes = [c_void_p, c_float]
        L.ref_total_inner_mass.argtypes = [c_void_p]
        L.ref_total_inner_mass.restype = c_float
        L.ref_add_source.argtypes = [c_int, _f32p, _f32p, c_float]
        L.ref_set_bnd.argtypes = [c_int, c_int, _f32p]
        L.ref_lin_solve.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.ref_diffuse.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.ref_diffuse_blur.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int]
        L.ref_advect.argtypes = [c_int, c_int, _f32p, _f32p, _f32p, _f32p, c_float]
        L.ref_advect_euler.argtypes = [c_int, c_int, _f32p, _f32p, _f32p, _f32p, c_float]
        L.ref_project.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_int]
        _u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
        L.ref_create_typed.restype = c_void_p
        L.ref_create_typed.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_int, c_int, c_float, c_int, c_int]
        L.ref_cell_info.restype = ctypes.POINTER(ctypes.c_ubyte)
        L.ref_cell_info.argtypes = [c_void_p]
        L.ref_make_cell_info.argtypes = [c_int, c_int, _u8p]
        L.ref_calc_border.argtypes = [c_int, _u8p]
        L.ref_lin_solve_b.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.ref_diffuse_blur_b.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.ref_advect_b.argtypes = [c_int, c_int, _f32p, _f32p, _f32p, _f32p, c_float]
        L.ref_project_b.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_int, _u8p]
        L.ref_lin_solve_bo.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, _u8p]
        L.ref_project_bo.argtypes = [c_int, _f32p, _f32p, _f32p, _f32p, c_int, _u8p]
        L.ref_lin_solve_bso.argtypes = [c_int, c_int, _f32p, _f32p, c_float, c_float, c_int, c_void_p, _f32p, c_int, _u8p]
        L.ref_p.restype = ctypes.POINTER(c_float)
        L.ref_p.argtypes = [c_void_p]
        L.ref_visc_air.restype = c_float
        L.ref_visc_air.argtypes = [c_void_p]

    class Sim:
        def __init__(self, ref, N, visc, diff, vorticity=0.0, iters=20, damp=0.0, fluids=None, fstype=0, square_border=False):
            """fluids: optional list of (diff, visc) per fluid (source field ids 2+k address fluid k).
            fstype: 0 ORIGINAL, 1 ORIGINAL_BORDERED, 2 ORIGINAL_BORDERED_MF (FluidSystem.h:10-17)."""
            self.ref, self.N = ref, N
            if fstype or square_border:
                fl = fluids or [(diff, visc)]
                d = np.array([f[0] for f in fl], np.float32)
                v = np.array([f[1] for f in fl], np.float32)
                self.h = ref.L.ref_create_typed(N, len(fl), d, v, vorticity, iters, 1 if vorticity != 0.0 else 0, damp, fstype,
                                                1 if square_border else 0)
            elif fluids:
                d = np.array([f[0] for f in fluids], np.float32)
                v = np.array([f[1] for f in fluids], np.float32)
                self.h = ref.L.ref_create_multi(N, len(fluids), d, v, vorticity, iters, 1 if vorticity != 0.0 else 0, damp)
            else:
                self.h = ref.L.ref_create(N, visc, diff, vorticity, iters, 1 if vorticity != 0.0 else 0, damp)

        def density(self, fluid, prev=False):
            return np.ctypeslib.as_array(self.ref.L.ref_density(self.h, fluid, 1 if prev else 0), ((self.N + 2) ** 2,))

        def set_sources(self, sources):
            n = len(sources)
            cols = [np.array([s[k] for s in sources], np.int32) for k in range(4)]
            val = np.array([s[4] for s in sources], np.float32)
            self.ref.L.ref_set_sources(self.h, n, cols[0].ctypes.data, cols[1].ctypes.data, cols[2].ctypes.data,
                                       cols[3].ctypes.data, val.ctypes.data)

        def field(self, which):
            """0=u 1=v 2=dens 3=u.Prev 4=v.Prev 5=dens.Prev -- a live view."""
            return np.ctypeslib.as_array(self.ref.L.ref_field(self.h, which), ((self.N + 2) ** 2,))

        def update(self, dt, nsteps=1):
            self.ref.L.ref_update(self.h, dt, nsteps)

        def vel_step(self, dt):
            self.ref.L.ref_vel_step(self.h, dt)

        def dens_step(self, dt):
            self.ref.L.ref_dens_step(self.h, dt)

        def total_inner_mass(self):
            return float(self.ref.L.ref_total_inner_mass(self.h))

        def p(self):
            return np.ctypeslib.as_array(self.ref.L.ref_p(self.h), ((self.N + 2) ** 2,))

        def cell_info(self):
            return np.ctypeslib.as_array(self.ref.L.ref_cell_info(self.h), ((self.N + 2) ** 2,))

        def close(self):
            if self.h:
                self.ref.L.ref_destroy(self.h)
                self.h = None

    def sim(self, N, visc, diff, vorticity=0.0, iters=20, damp=0.0, fluids=None, fstype=0, square_border=False):
        return Ref.Sim(self, N, visc, diff, vorticity, iters, damp, fluids, fstype, square_border)
