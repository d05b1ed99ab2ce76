"""ctypes view of libfluid_cuda.so (include/fluid_cuda.h) for the tests and bench.py.

The product is the C-ABI shared library; host applications are C (see
host/fluid_host_demo.c and INTEGRATION.md).  This module adds no arithmetic and
no fallback: if the library is missing, or there is no CUDA device, calls fail
loudly (``FluidError``).

``FluidSystem`` mirrors the names of the reference class it stands in for
(FluidSystem.h:20-69: Setup / Update / GetVelX / GetVelY / GetDensities / Clear)
so that parity tests read like reference code.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libfluid_cuda.so")

GS_WAVEFRONT, GS_REDBLACK = 0, 1
ADVECT_MIDPOINT, ADVECT_EULER = 0, 1
U, V, DENS, U_PREV, V_PREV, DENS_PREV = range(6)
SRC_PREV_SET, SRC_CURR_ADD, SRC_CURR_SET = 0, 1, 2
TYPE_ORIGINAL, TYPE_ORIGINAL_BORDERED, TYPE_ORIGINAL_BORDERED_MF = 0, 1, 2
K_LIN_SOLVE, K_BLUR, K_ADVECT, K_ADVECT_VEL, K_PROJECT, K_ADD_SOURCE = range(6)

# every symbol include/fluid_cuda.h declares (tests check the .so exports them all)
EXPORTS = [
    "fluid_last_error", "fluid_version", "fluid_launch_count", "fluid_launch_count_reset",
    "fluid_set_gs_order", "fluid_set_gs_iters", "fluid_set_advect_mode", "fluid_set_scratch_download",
    "fluid_set_device", "fluid_host_alloc", "fluid_host_free",
    "fluid_add_source", "fluid_set_bnd", "fluid_lin_solve", "fluid_diffuse", "fluid_diffuse_blur", "fluid_advect",
    "fluid_project", "fluid_vel_step", "fluid_dens_step", "fluid_update", "fluid_vort_confinement", "fluid_damp",
    "fluid_params_default", "fluid_ctx_create", "fluid_ctx_destroy", "fluid_ctx_clear", "fluid_ctx_upload",
    "fluid_ctx_download", "fluid_ctx_set_sources", "fluid_ctx_vel_step", "fluid_ctx_dens_step", "fluid_ctx_step",
    "fluid_ctx_sync", "fluid_ctx_last_step_ms", "fluid_ctx_set_modes", "fluid_ctx_set_coeffs", "fluid_ctx_set_fluid",
    "fluid_ctx_set_profiling", "fluid_ctx_profile", "fluid_ctx_set_tuning",
    "fluid_ctx_set_fast_division", "fluid_ctx_division_window", "fluid_ctx_set_small_grid_kernels",
    "fluid_ctx_set_wavefront_tile", "fluid_ctx_set_wavefront_persistent", "fluid_ctx_render", "fluid_ctx_set_stream_kernel", "fluid_ctx_set_overlap", "fluid_ctx_set_advect_kernel", "fluid_ctx_set_sparse_sources",
    "fluid_ctx_reset_cell_info", "fluid_ctx_disable_edges", "fluid_ctx_add_square_border", "fluid_ctx_set_cell_info",
    "fluid_ctx_get_cell_info", "fluid_ctx_set_visc_air",
    "fluid_ctx_peer_halo_local", "fluid_ctx_peer_halo_connect", "fluid_ipc_export", "fluid_ipc_import", "fluid_ipc_close",
    "fluid_ctx_set_exchange", "fluid_ctx_halo_refresh", "fluid_ctx_set_stream", "fluid_ctx_halo_regions", "fluid_ctx_upload_local", "fluid_ctx_download_local",
    "fluid_ctx_local_rows",
    "fluid_ctx_add_source", "fluid_ctx_set_bnd", "fluid_ctx_lin_solve", "fluid_ctx_blur", "fluid_ctx_advect",
    "fluid_ctx_advect_velocity", "fluid_ctx_project", "fluid_ctx_time_kernel",
    "fluid_lin_solve_rb_static",
    "fluid_multi_create", "fluid_multi_destroy", "fluid_multi_count", "fluid_multi_set_sources", "fluid_multi_upload",
    "fluid_multi_download", "fluid_multi_step", "fluid_multi_sync", "fluid_multi_last_step_ms",
]


class FluidError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("fluid_cuda error %d: %s" % (code, msg))
        self.code = code


class Params(ctypes.Structure):
    _fields_ = [("N", ctypes.c_int), ("batch", ctypes.c_int), ("gs_iters", ctypes.c_int), ("gs_order", ctypes.c_int),
                ("advect_mode", ctypes.c_int), ("device", ctypes.c_int), ("visc", ctypes.c_float),
                ("diff", ctypes.c_float), ("vorticity", ctypes.c_float), ("damp", ctypes.c_float),
                ("slab_index", ctypes.c_int), ("slab_count", ctypes.c_int), ("slab_rows", ctypes.c_int),
                ("slab_halo", ctypes.c_int), ("n_fluids", ctypes.c_int), ("fs_type", ctypes.c_int)]


class Source(ctypes.Structure):
    _fields_ = [("field", ctypes.c_int), ("i", ctypes.c_int), ("j", ctypes.c_int), ("mode", ctypes.c_int),
                ("value", ctypes.c_float)]


_f32 = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_ci, _cf, _vp = ctypes.c_int, ctypes.c_float, ctypes.c_void_p

EXCHANGE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int))

_lib = None


def lib():
    """Load libfluid_cuda.so once; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FluidError(-1, "%s not built (run `python -c 'import __graft_entry__ as g; g.build()'` or "
                             "`make -C fluid-simulation_b200`)" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    L.fluid_last_error.restype = ctypes.c_char_p
    L.fluid_version.restype = ctypes.c_char_p
    L.fluid_launch_count.restype = ctypes.c_ulonglong
    L.fluid_launch_count_reset.restype = None
    L.fluid_host_alloc.argtypes = [ctypes.POINTER(_vp), ctypes.c_size_t]
    L.fluid_host_free.argtypes = [_vp]
    L.fluid_add_source.argtypes = [_ci, _f32, _f32, _cf]
    L.fluid_set_bnd.argtypes = [_ci, _ci, _f32]
    L.fluid_lin_solve.argtypes = [_ci, _ci, _f32, _f32, _cf, _cf]
    L.fluid_diffuse.argtypes = [_ci, _ci, _f32, _f32, _cf, _cf]
    L.fluid_diffuse_blur.argtypes = [_ci, _ci, _f32, _f32, _cf, _cf]
    L.fluid_advect.argtypes = [_ci, _ci, _f32, _f32, _f32, _f32, _cf]
    L.fluid_project.argtypes = [_ci, _f32, _f32, _f32, _f32]
    L.fluid_vel_step.argtypes = [_ci, _f32, _f32, _f32, _f32, _cf, _cf]
    L.fluid_dens_step.argtypes = [_ci, _f32, _f32, _f32, _f32, _cf, _cf]
    L.fluid_update.argtypes = [_ci, _f32, _f32, _f32, _ci, _vp, _cf, _cf, _cf, _cf, _cf]
    L.fluid_vort_confinement.argtypes = [_ci, _f32, _f32, _cf]
    L.fluid_damp.argtypes = [_ci, _f32, _cf, _cf]
    L.fluid_params_default.argtypes = [ctypes.POINTER(Params)]
    L.fluid_params_default.restype = None
    L.fluid_ctx_create.argtypes = [ctypes.POINTER(Params), ctypes.POINTER(_vp)]
    L.fluid_ctx_destroy.argtypes = [_vp]
    L.fluid_ctx_clear.argtypes = [_vp]
    L.fluid_ctx_upload.argtypes = [_vp, _ci, _ci, _f32]
    L.fluid_ctx_download.argtypes = [_vp, _ci, _ci, _f32]
    L.fluid_ctx_set_sources.argtypes = [_vp, _ci, _ci, _vp]
    L.fluid_ctx_vel_step.argtypes = [_vp, _cf]
    L.fluid_ctx_dens_step.argtypes = [_vp, _cf]
    L.fluid_ctx_step.argtypes = [_vp, _cf, _ci]
    L.fluid_ctx_sync.argtypes = [_vp]
    L.fluid_ctx_last_step_ms.argtypes = [_vp, ctypes.POINTER(_cf)]
    L.fluid_ctx_set_modes.argtypes = [_vp, _ci, _ci, _ci]
    L.fluid_ctx_set_coeffs.argtypes = [_vp, _cf, _cf, _cf, _cf]
    L.fluid_ctx_set_fluid.argtypes = [_vp, _ci, _cf, _cf]
    L.fluid_ctx_set_tuning.argtypes = [_vp, _ci]
    L.fluid_ctx_set_fast_division.argtypes = [_vp, _ci]
    L.fluid_ctx_set_small_grid_kernels.argtypes = [_vp, _ci]
    L.fluid_ctx_set_wavefront_tile.argtypes = [_vp, _ci]
    L.fluid_ctx_set_wavefront_persistent.argtypes = [_vp, _ci]
    L.fluid_ctx_set_stream_kernel.argtypes = [_vp, _ci]
    L.fluid_ctx_set_overlap.argtypes = [_vp, _ci]
    L.fluid_ctx_set_advect_kernel.argtypes = [_vp, _ci]
    L.fluid_ctx_set_sparse_sources.argtypes = [_vp, _ci]
    _u8 = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
    L.fluid_ctx_reset_cell_info.argtypes = [_vp]
    L.fluid_ctx_disable_edges.argtypes = [_vp, _ci, ctypes.POINTER(_ci)]
    L.fluid_ctx_add_square_border.argtypes = [_vp]
    L.fluid_ctx_set_cell_info.argtypes = [_vp, _u8, _ci]
    L.fluid_ctx_get_cell_info.argtypes = [_vp, _u8]
    L.fluid_ctx_set_visc_air.argtypes = [_vp, _cf]
    L.fluid_ctx_peer_halo_local.argtypes = [_vp, ctypes.POINTER(_vp), ctypes.POINTER(ctypes.c_size_t)]
    L.fluid_ctx_peer_halo_connect.argtypes = [_vp, _vp, _vp]
    L.fluid_ipc_export.argtypes = [_vp, ctypes.c_char_p]
    L.fluid_ipc_import.argtypes = [ctypes.c_char_p, ctypes.POINTER(_vp)]
    L.fluid_ipc_close.argtypes = [_vp]
    L.fluid_ctx_render.argtypes = [_vp, _ci, _f32, _f32, _vp]
    L.fluid_ctx_division_window.argtypes = [_vp, _cf, ctypes.POINTER(_ci), ctypes.POINTER(_ci)]
    L.fluid_ctx_upload_local.argtypes = [_vp, _ci, _f32]
    L.fluid_ctx_download_local.argtypes = [_vp, _ci, _f32]
    L.fluid_ctx_local_rows.argtypes = [_vp, ctypes.POINTER(_ci), ctypes.POINTER(_ci)]
    L.fluid_ctx_set_exchange.argtypes = [_vp, EXCHANGE_FN, _vp]
    L.fluid_ctx_halo_refresh.argtypes = [_vp, _ci, ctypes.POINTER(_ci)]
    L.fluid_ctx_set_stream.argtypes = [_vp, _vp, _ci]
    L.fluid_ctx_halo_regions.argtypes = [_vp, _ci] + [ctypes.POINTER(_vp)] * 4 + [ctypes.POINTER(ctypes.c_size_t)]
    L.fluid_ctx_set_profiling.argtypes = [_vp, _ci]
    L.fluid_ctx_profile.argtypes = [_vp, ctypes.POINTER(_cf), ctypes.POINTER(ctypes.c_ulonglong), _ci]
    L.fluid_ctx_add_source.argtypes = [_vp, _ci, _ci, _cf]
    L.fluid_ctx_set_bnd.argtypes = [_vp, _ci, _ci]
    L.fluid_ctx_lin_solve.argtypes = [_vp, _ci, _ci, _ci, _cf, _cf]
    L.fluid_ctx_blur.argtypes = [_vp, _ci, _ci, _ci, _cf, _cf]
    L.fluid_ctx_advect.argtypes = [_vp, _ci, _ci, _ci, _ci, _ci, _cf]
    L.fluid_ctx_advect_velocity.argtypes = [_vp, _ci, _ci, _ci, _ci, _cf]
    L.fluid_ctx_project.argtypes = [_vp, _ci, _ci, _ci, _ci]
    L.fluid_ctx_time_kernel.argtypes = [_vp, _ci, _ci, ctypes.POINTER(_cf)]
    L.fluid_lin_solve_rb_static.argtypes = [_ci, _ci, _vp, _vp, _cf, _cf, _ci, _vp]
    L.fluid_multi_create.argtypes = [ctypes.POINTER(Params), _ci, ctypes.POINTER(_ci), ctypes.POINTER(_vp)]
    L.fluid_multi_destroy.argtypes = [_vp]
    L.fluid_multi_count.argtypes = [_vp]
    L.fluid_multi_set_sources.argtypes = [_vp, _ci, _vp]
    L.fluid_multi_upload.argtypes = [_vp, _ci, _f32]
    L.fluid_multi_download.argtypes = [_vp, _ci, _f32]
    L.fluid_multi_step.argtypes = [_vp, _cf, _ci]
    L.fluid_multi_sync.argtypes = [_vp]
    L.fluid_multi_last_step_ms.argtypes = [_vp, ctypes.POINTER(_cf)]
    _lib = L
    return L


def dens_of(k):
    """Field id of the density of fluid k (FLUID_DENS_OF in the header)."""
    return DENS if k == 0 else 4 + 2 * k


def dens_prev_of(k):
    return DENS_PREV if k == 0 else 5 + 2 * k


def check(rc):
    if rc != 0:
        raise FluidError(rc, lib().fluid_last_error().decode())


def set_modes(gs_order=GS_WAVEFRONT, gs_iters=20, advect_mode=ADVECT_MIDPOINT, scratch_download=True):
    """Modes of the Stam-signature host-pointer calls (thread-local in the library)."""
    L = lib()
    check(L.fluid_set_gs_order(gs_order))
    check(L.fluid_set_gs_iters(gs_iters))
    check(L.fluid_set_advect_mode(advect_mode))
    check(L.fluid_set_scratch_download(1 if scratch_download else 0))


# Stam-signature primitives on host numpy arrays (flat float32, (N+2)^2, in place)
def add_source(N, x, s, dt): check(lib().fluid_add_source(N, x, s, dt))
def set_bnd(N, b, x): check(lib().fluid_set_bnd(N, b, x))
def lin_solve(N, b, x, x0, a, c): check(lib().fluid_lin_solve(N, b, x, x0, a, c))
def diffuse(N, b, x, x0, diff, dt): check(lib().fluid_diffuse(N, b, x, x0, diff, dt))
def diffuse_blur(N, b, x, x0, diff, dt): check(lib().fluid_diffuse_blur(N, b, x, x0, diff, dt))
def advect(N, b, d, d0, u, v, dt): check(lib().fluid_advect(N, b, d, d0, u, v, dt))
def project(N, u, v, p, div): check(lib().fluid_project(N, u, v, p, div))
def vel_step(N, u, v, u0, v0, visc, dt): check(lib().fluid_vel_step(N, u, v, u0, v0, visc, dt))
def dens_step(N, x, x0, u, v, diff, dt): check(lib().fluid_dens_step(N, x, x0, u, v, diff, dt))
def update(N, u, v, dens, sources, visc, diff, dt, vorticity=0.0, damp=0.0):
    """FluidSystem::Update on host arrays; sources = [(field, i, j, mode, value), ...]."""
    sources = list(sources)
    arr = (Source * max(1, len(sources)))(*[Source(*s) for s in sources])
    check(lib().fluid_update(N, u, v, dens, len(sources), ctypes.cast(arr, _vp), visc, diff, vorticity, damp, dt))


def vort_confinement(N, u, v, vorticity): check(lib().fluid_vort_confinement(N, u, v, vorticity))
def damp(N, vel, amount, dt): check(lib().fluid_damp(N, vel, amount, dt))


def ipc_export(device_ptr):
    """64-byte CUDA IPC handle of a device allocation of this process."""
    buf = ctypes.create_string_buffer(64)
    check(lib().fluid_ipc_export(device_ptr, buf))
    return buf.raw


def ipc_import(handle):
    """Device address, in this process, of an allocation another process exported."""
    p = _vp()
    check(lib().fluid_ipc_import(handle, ctypes.byref(p)))
    return p.value


def ipc_close(device_ptr): check(lib().fluid_ipc_close(device_ptr))


def pinned_empty(n_floats):
    """float32 numpy array backed by pinned host memory from fluid_host_alloc."""
    p = _vp()
    check(lib().fluid_host_alloc(ctypes.byref(p), n_floats * 4))
    arr = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_float)), (n_floats,))
    return arr, p


class FluidSystem:
    """Resident GPU context with the reference class's method names
    (FluidSystem.h:20-69; FSType::ORIGINAL, one fluid)."""

    def __init__(self):
        self.h = None
        self.N = 0
        self.batch = 1

    def Setup(self, N, visc=0.0, diff=0.0, vorticity=0.0, damp=0.0, gs_order=GS_WAVEFRONT, gs_iters=20,
              advect_mode=ADVECT_MIDPOINT, batch=1, device=-1, slab=None, slab_halo=0, n_fluids=1, fs_type=TYPE_ORIGINAL):
        """slab = (index, count, rows): this context is one row slab of a global N x (rows*count) grid;
        slab = (0, 1, rows): a whole non-square grid of N x rows."""
        self.close()
        p = Params()
        lib().fluid_params_default(ctypes.byref(p))
        p.N, p.batch, p.gs_iters, p.gs_order, p.advect_mode, p.device = N, batch, gs_iters, gs_order, advect_mode, device
        p.visc, p.diff, p.vorticity, p.damp = visc, diff, vorticity, damp
        p.n_fluids = n_fluids
        p.fs_type = fs_type
        self.NYG = N
        if slab is not None:
            p.slab_index, p.slab_count, p.slab_rows = slab
            p.slab_halo = slab_halo
            self.NYG = slab[1] * slab[2]
        h = _vp()
        check(lib().fluid_ctx_create(ctypes.byref(p), ctypes.byref(h)))
        self.h, self.N, self.batch = h, N, batch
        self._exchange_cb = None
        return self

    def SetExchange(self, fn):
        """fn(field_ids: list[int]) -> None is called at every halo exchange point (row slabs)."""
        def thunk(_user, n, ids):
            try:
                fn([ids[k] for k in range(n)])
                return 0
            except Exception:  # never let an exception cross the C boundary
                import traceback
                traceback.print_exc()
                return 1
        self._exchange_cb = EXCHANGE_FN(thunk)  # keep a reference alive
        check(lib().fluid_ctx_set_exchange(self.h, self._exchange_cb, None))

    def HaloRefresh(self, fields):
        arr = (ctypes.c_int * len(fields))(*fields)
        check(lib().fluid_ctx_halo_refresh(self.h, len(fields), arr))

    def LocalRows(self):
        """(rows stored incl. halo and border rows, global row index of the first)."""
        n, j0 = _ci(), _ci()
        check(lib().fluid_ctx_local_rows(self.h, ctypes.byref(n), ctypes.byref(j0)))
        return n.value, j0.value

    def upload_local(self, field, host): check(lib().fluid_ctx_upload_local(self.h, field, host))
    def download_local(self, field, host): check(lib().fluid_ctx_download_local(self.h, field, host))

    def PeerHaloLocal(self):
        """(device address, bytes) of this slab's inbox/flag block (allocated on first call)."""
        p, n = _vp(), ctypes.c_size_t()
        check(lib().fluid_ctx_peer_halo_local(self.h, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def PeerHaloConnect(self, lower_block, upper_block):
        check(lib().fluid_ctx_peer_halo_connect(self.h, lower_block, upper_block))

    def SetStream(self, stream_handle, exchange_stream_ordered=False):
        check(lib().fluid_ctx_set_stream(self.h, stream_handle, 1 if exchange_stream_ordered else 0))

    def HaloRegions(self, field):
        """(send_down, recv_down, send_up, recv_up, count_floats) device addresses (None where no neighbour)."""
        p = [_vp() for _ in range(4)]
        n = ctypes.c_size_t()
        check(lib().fluid_ctx_halo_regions(self.h, field, *[ctypes.byref(q) for q in p], ctypes.byref(n)))
        return tuple(q.value for q in p) + (n.value,)

    def close(self):
        if self.h:
            lib().fluid_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def Clear(self): check(lib().fluid_ctx_clear(self.h))

    def SetSources(self, sources, sim=-1):
        """sources: iterable of (field, i, j, mode, value)."""
        sources = list(sources)
        arr = (Source * max(1, len(sources)))(*[Source(*s) for s in sources])
        check(lib().fluid_ctx_set_sources(self.h, sim, len(sources), ctypes.cast(arr, _vp)))

    def Update(self, dt, nsteps=1):
        check(lib().fluid_ctx_step(self.h, dt, nsteps))

    def Sync(self): check(lib().fluid_ctx_sync(self.h))

    def LastStepMs(self):
        ms = _cf()
        check(lib().fluid_ctx_last_step_ms(self.h, ctypes.byref(ms)))
        return ms.value

    def VelStep(self, dt): check(lib().fluid_ctx_vel_step(self.h, dt))
    def DensStep(self, dt): check(lib().fluid_ctx_dens_step(self.h, dt))

    def SetModes(self, gs_order, gs_iters=20, advect_mode=ADVECT_MIDPOINT):
        check(lib().fluid_ctx_set_modes(self.h, gs_order, gs_iters, advect_mode))

    def SetFluid(self, k, diff, visc):
        """FluidProps of fluid k (its viscosity enters through visc_avg, FluidSystem.cpp:26-30)."""
        check(lib().fluid_ctx_set_fluid(self.h, k, diff, visc))

    def SetCoeffs(self, visc, diff, vorticity=0.0, damp=0.0):
        check(lib().fluid_ctx_set_coeffs(self.h, visc, diff, vorticity, damp))

    def upload(self, field, host, sim=0):
        check(lib().fluid_ctx_upload(self.h, field, sim, np.ascontiguousarray(host, np.float32).ravel()))

    def download(self, field, sim=0, out=None):
        if out is None:
            out = np.empty((self.N + 2) * (self.NYG + 2), np.float32)
        check(lib().fluid_ctx_download(self.h, field, sim, out))
        return out

    def GetVelX(self, sim=0): return self.download(U, sim)
    def GetVelY(self, sim=0): return self.download(V, sim)
    def GetDensities(self, sim=0): return self.download(DENS, sim)

    def SetTuning(self, rb_tile_sweeps): check(lib().fluid_ctx_set_tuning(self.h, rb_tile_sweeps))

    def Render(self, color=(1.0, 1.0, 1.0), sim=0, lines=True):
        """FluidRenderer::DoBuffer headless: (pixels[(N+2)^2,3], lines[(N+2)^2,4] or None).
        color: 3 floats per fluid of the context."""
        n = (self.N + 2) ** 2
        px = np.empty(n * 3, np.float32)
        ln = np.empty(n * 4, np.float32) if lines else None
        col = np.ascontiguousarray(np.asarray(color, np.float32).ravel())
        check(lib().fluid_ctx_render(self.h, sim, col, px, ln.ctypes.data if lines else None))
        return px.reshape(n, 3), (ln.reshape(n, 4) if lines else None)

    def SetStreamKernel(self, on): check(lib().fluid_ctx_set_stream_kernel(self.h, (1 if on else 0) if isinstance(on, bool) else int(on)))
    def SetOverlap(self, on): check(lib().fluid_ctx_set_overlap(self.h, 1 if on else 0))
    def SetAdvectKernel(self, which): check(lib().fluid_ctx_set_advect_kernel(self.h, which))
    def SetSparseSources(self, on): check(lib().fluid_ctx_set_sparse_sources(self.h, 1 if on else 0))

    # cell map of a TYPE_ORIGINAL_BORDERED context (FluidSystem::GetInfo / DisableEdge / AddBorder(SquareBorder))
    def SetViscAir(self, v): check(lib().fluid_ctx_set_visc_air(self.h, v))
    def ResetCellInfo(self): check(lib().fluid_ctx_reset_cell_info(self.h))
    def AddSquareBorder(self): check(lib().fluid_ctx_add_square_border(self.h))

    def DisableEdges(self, edges):
        flat = [int(v) for e in edges for v in e]
        arr = (ctypes.c_int * max(1, len(flat)))(*flat)
        check(lib().fluid_ctx_disable_edges(self.h, len(flat) // 4, arr))

    def SetCellInfo(self, info, recalc=False):
        check(lib().fluid_ctx_set_cell_info(self.h, np.ascontiguousarray(info, np.uint8).ravel(), 1 if recalc else 0))

    def GetInfo(self):
        out = np.empty((self.N + 2) ** 2, np.uint8)
        check(lib().fluid_ctx_get_cell_info(self.h, out))
        return out

    def SetWavefrontTile(self, tile): check(lib().fluid_ctx_set_wavefront_tile(self.h, tile))
    def SetWavefrontPersistent(self, on):
        """0 / False = one launch per hyperplane of tiles, 1 = automatic (the default: the row pipeline), 2 / True = one launch per
        solve with tile tickets, 3 = the row pipeline (kernels_rowwave.cuh)."""
        check(lib().fluid_ctx_set_wavefront_persistent(self.h, (2 if on else 0) if isinstance(on, bool) else int(on)))

    def SetSmallGridKernels(self, on): check(lib().fluid_ctx_set_small_grid_kernels(self.h, 1 if on else 0))

    def SetFastDivision(self, mode): check(lib().fluid_ctx_set_fast_division(self.h, mode))

    def DivisionWindow(self, divisor):
        lo, hi = _ci(), _ci()
        check(lib().fluid_ctx_division_window(self.h, divisor, ctypes.byref(lo), ctypes.byref(hi)))
        return lo.value, hi.value

    PROF_CLASSES = ("lin_solve", "blur", "advect", "project_aux", "add_source", "other", "exchange")

    def SetProfiling(self, on=True): check(lib().fluid_ctx_set_profiling(self.h, 1 if on else 0))

    def Profile(self, reset=True):
        """{class: (ms, launches)} accumulated since the last reset."""
        ms = (_cf * 7)()
        n = (ctypes.c_ulonglong * 7)()
        check(lib().fluid_ctx_profile(self.h, ms, n, 1 if reset else 0))
        return {k: (ms[i], int(n[i])) for i, k in enumerate(self.PROF_CLASSES)}

    # single primitives on resident fields
    def add_source(self, x, s, dt): check(lib().fluid_ctx_add_source(self.h, x, s, dt))
    def set_bnd(self, b, x): check(lib().fluid_ctx_set_bnd(self.h, b, x))
    def lin_solve(self, b, x, x0, a, c): check(lib().fluid_ctx_lin_solve(self.h, b, x, x0, a, c))
    def blur(self, b, x, x0, a, c): check(lib().fluid_ctx_blur(self.h, b, x, x0, a, c))
    def advect(self, b, d, d0, u, v, dt): check(lib().fluid_ctx_advect(self.h, b, d, d0, u, v, dt))
    def advect_velocity(self, u, v, u0, v0, dt): check(lib().fluid_ctx_advect_velocity(self.h, u, v, u0, v0, dt))
    def project(self, u, v, p, div): check(lib().fluid_ctx_project(self.h, u, v, p, div))

    def time_kernel(self, which, reps=5):
        ms = _cf()
        check(lib().fluid_ctx_time_kernel(self.h, which, reps, ctypes.byref(ms)))
        return ms.value


class FluidMulti:
    """fluid_multi_*: several GPUs (or several slabs on one GPU) from one process, peer-copy halos."""

    def __init__(self, N, rows, devices, visc=0.0, diff=0.0, gs_iters=20, advect_mode=ADVECT_MIDPOINT, slab_halo=0):
        self.h = None
        p = Params()
        lib().fluid_params_default(ctypes.byref(p))
        p.N, p.gs_iters, p.gs_order, p.advect_mode = N, gs_iters, GS_REDBLACK, advect_mode
        p.visc, p.diff, p.slab_rows, p.slab_halo = visc, diff, rows, slab_halo
        dev = (ctypes.c_int * len(devices))(*devices)
        h = _vp()
        check(lib().fluid_multi_create(ctypes.byref(p), len(devices), dev, ctypes.byref(h)))
        self.h, self.N, self.NYG = h, N, rows * len(devices)

    def SetSources(self, sources):
        sources = list(sources)
        arr = (Source * max(1, len(sources)))(*[Source(*s) for s in sources])
        check(lib().fluid_multi_set_sources(self.h, len(sources), ctypes.cast(arr, _vp)))

    def upload(self, field, host_global): check(lib().fluid_multi_upload(self.h, field, host_global))

    def download(self, field):
        out = np.zeros((self.N + 2) * (self.NYG + 2), np.float32)
        check(lib().fluid_multi_download(self.h, field, out))
        return out

    def Update(self, dt, nsteps=1): check(lib().fluid_multi_step(self.h, dt, nsteps))
    def Sync(self): check(lib().fluid_multi_sync(self.h))

    def LastStepMs(self):
        ms = _cf()
        check(lib().fluid_multi_last_step_ms(self.h, ctypes.byref(ms)))
        return ms.value

    def close(self):
        if self.h:
            lib().fluid_multi_destroy(self.h)
            self.h = None

    __del__ = close
