/*
 * fluid_cuda.h -- C ABI of libfluid_cuda.so: the Stam "Stable Fluids" step of
 * surajiyer/Fluid-Simulation (FSType::ORIGINAL pipeline) as hand-written
 * sm_100a CUDA kernels.  Pure C, plain pointers and sizes, no C++/torch types.
 *
 * Reference citations are file:line into /root/reference/FluidSim/src/.
 *
 * Grids.  A host field is a dense (N+2)*(N+2) float array, cell (i,j) at
 * i + (N+2)*j with i contiguous (FluidSystem::ID, FluidSystem.h:63-65); i,j in
 * 1..N are interior cells, 0 and N+1 the border.  Device storage is private
 * (row-pitched) and only crosses this boundary by copy.
 *
 * Numerics.  The library is compiled with -fmad=false -prec-div=true
 * -prec-sqrt=true -ftz=false and evaluates every expression with the
 * reference's association, so with FLUID_GS_WAVEFRONT every entry point is
 * bit-identical to the reference CPU solver; FLUID_GS_REDBLACK changes only
 * the relaxation order (and is bit-identical to a CPU red-black statement).
 *
 * Errors.  Every function returns 0 (FLUID_OK) or a fluid_status code and
 * never aborts; fluid_last_error() gives a thread-local message.  There is no
 * CPU fallback: without a usable CUDA device every call fails with
 * FLUID_ERR_CUDA.  (The reference has no error handling at all: every
 * primitive is void, FluidSystem.h:71-140.)
 *
 * Threading.  One context = one host thread at a time.  The host-pointer
 * entry points share one internal context per thread.
 */
#ifndef FLUID_CUDA_H
#define FLUID_CUDA_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum fluid_status {
	FLUID_OK = 0,
	FLUID_ERR_INVALID = 1,     /* bad argument */
	FLUID_ERR_CUDA = 2,        /* CUDA runtime/driver error, or no device */
	FLUID_ERR_NOMEM = 3,       /* device or host allocation failed */
	FLUID_ERR_UNSUPPORTED = 4  /* valid request this build cannot serve */
} fluid_status;

/* Relaxation order for lin_solve / diffuse / project. */
typedef enum fluid_gs_order {
	FLUID_GS_WAVEFRONT = 0, /* reference order (lexicographic, in place), bitwise == CPU */
	FLUID_GS_REDBLACK = 1   /* fast path: red (i+j even), black, set_bnd per sweep */
} fluid_gs_order;

/* Backtrace used by advect. */
typedef enum fluid_advect_mode {
	FLUID_ADVECT_MIDPOINT = 0, /* MidpointBackward, FluidSystem.h:185-206 (what Advect1 calls) */
	FLUID_ADVECT_EULER = 1     /* EulerBack, FluidSystem.h:208-219 (Stam's original, uncalled) */
} fluid_advect_mode;

/* Solver type (FluidSystem::FSType, FluidSystem.h:10-17; SetType, FluidSystem.cpp:69-99).
 * ORIGINAL          = VelStep1  + DensStep1   (the path north_star names; every fast kernel serves it)
 * ORIGINAL_BORDERED = VelStep1B + DensStep1BS (FluidSystem.cpp:378-398, 301-311): the same stencils masked per cell by
 *                     cellInfo (FluidCode.h:61-73) -- internal walls, object cells -- with no SetBorder in solve,
 *                     projection and advection.  Its first AddArrDt adds the Y source to vX, as the reference does
 *                     (FluidSystem.cpp:380).  Every primitive of such a context (lin_solve, blur, project, advect,
 *                     vel_step, dens_step, step) is the B variant; whole grids only (no row slabs). */
typedef enum fluid_fs_type {
	FLUID_TYPE_ORIGINAL = 0,
	FLUID_TYPE_ORIGINAL_BORDERED = 1,
	/* VelStep1BSO + DensStep1BS (FluidSystem.cpp:424-447, 301-311): what the shipped scene runs (SystemManager.cpp:108).
	 * The velocity diffuses with a viscosity mixed per cell from the densities of all fluids (LinearSolve1BSO, :673-720;
	 * cells without fluid take fluid_ctx_set_visc_air, default 0.001 as FluidSystem.h:149) and the projection is ProjectBO
	 * (:1061-1122), which solves into the system's own pressure array and lets the velocity stored in non-FLUID neighbour
	 * cells enter the divergence and the solve (LinearSolve1BO, :580-615).  The reference's quirks are reproduced: the
	 * Y source added to vX (:426), bScX used for all four sides in LinearSolve1BO (:605-608), p_old ignored (:1096-1098).
	 * Rigid colliders themselves are out of scope: the cell map is static until the host changes it. */
	FLUID_TYPE_ORIGINAL_BORDERED_MF = 2
} fluid_fs_type;

/* cellInfo bits (enum CellInfo, FluidCode.h:61-73) */
#define FLUID_CELL_LEFT 1
#define FLUID_CELL_RIGHT 2
#define FLUID_CELL_TOP 4
#define FLUID_CELL_BOTTOM 8
#define FLUID_CELL_FLUID 16
#define FLUID_CELL_OBJECT 32

/* Field ids of a context: the Curr/Prev halves of the reference's FlipFlopArr
 * members vX, vY, densList[0] (FluidSystem.h:157-160, FluidCode.h:15-44). */
typedef enum fluid_field {
	FLUID_U = 0, FLUID_V = 1, FLUID_DENS = 2,
	FLUID_U_PREV = 3, FLUID_V_PREV = 4, FLUID_DENS_PREV = 5,
	FLUID_FIELD_COUNT = 6
} fluid_field;
/* More than one fluid (fluid_params.n_fluids, the reference's densList, FluidSystem.h:160): the
 * Curr / Prev density fields of fluid k >= 1 have the ids below; fluid 0 keeps FLUID_DENS(_PREV). */
#define FLUID_MAX_FLUIDS 8
#define FLUID_DENS_OF(k) ((k) == 0 ? (int) FLUID_DENS : 4 + 2 * (k))
#define FLUID_DENS_PREV_OF(k) ((k) == 0 ? (int) FLUID_DENS_PREV : 5 + 2 * (k))

/* How a forcing entry is applied each step (between the zero fill of the Prev
 * arrays and the velocity step, i.e. where FluidSystem::Update calls its
 * updaters, FluidSystem.cpp:104-118). */
typedef enum fluid_source_mode {
	FLUID_SRC_PREV_SET = 0, /* Prev[i,j]  = value  (Stam-style source, scaled by dt in add_source) */
	FLUID_SRC_CURR_ADD = 1, /* Curr[i,j] += value  (AddFluid density, FluidSystem.cpp:250-257) */
	FLUID_SRC_CURR_SET = 2  /* Curr[i,j]  = value  (SetVelocity, FluidSystem.cpp:135-140) */
} fluid_source_mode;

typedef struct fluid_source {
	int field; /* FLUID_U, FLUID_V or FLUID_DENS */
	int i, j;  /* cell, 0..N+1 */
	int mode;  /* fluid_source_mode */
	float value;
} fluid_source;

/* What FluidSystem::Setup fixes (FluidSystem.cpp:18-67) plus the knobs the
 * reference hard-codes.  Zero-initialise, then fluid_params_default(). */
typedef struct fluid_params {
	int N;            /* interior cells per side */
	int batch;        /* independent simulations advanced together (ensemble), >= 1 */
	int gs_iters;     /* FluidSystem::steps, 20 (FluidSystem.cpp:24) */
	int gs_order;     /* fluid_gs_order */
	int advect_mode;  /* fluid_advect_mode */
	int device;       /* CUDA device ordinal; -1 = current device */
	float visc;       /* visc_avg (FluidSystem.cpp:26-30) */
	float diff;       /* FluidProps::Diff of the one fluid */
	float vorticity;  /* 0 = confinement off (useVort=false) */
	float damp;       /* 0 = no damping (reference default 0.1, FluidSystem.h:150) */
	/* Row slab of a taller grid (multi-GPU): this context stores rows
	 * slab_index*slab_rows+1 .. (slab_index+1)*slab_rows of a global N x (slab_rows*slab_count)
	 * grid, plus slab_halo halo rows towards each neighbour.  slab_count <= 1: whole grid -- square
	 * (slab_rows = 0) or N columns x slab_rows rows (the reference is square-only; the scale factors
	 * dt*N, /N, 0.5*N, dt*diff*N*N keep using N, i.e. h = 1/N). */
	int slab_index, slab_count, slab_rows;
	int slab_halo;    /* 0 = default 48 rows; >= 44 */
	int n_fluids;     /* density fields advanced by dens_step (fluid_count, FluidSystem.cpp:21); 0 = 1 */
	int fs_type;      /* fluid_fs_type: which pair of step functions SetType would pick (FluidSystem.cpp:69-99); 0 = ORIGINAL */
} fluid_params;

typedef struct fluid_ctx fluid_ctx;

/* ---- library ----------------------------------------------------------- */

/* Thread-local description of the last failure ("" if none). */
const char *fluid_last_error(void);
/* "major.minor sm_100a ..." build string. */
const char *fluid_version(void);
/* Number of this library's kernels launched by the calling thread since the
 * last reset (bench.py reports it as gpu_launches). */
unsigned long long fluid_launch_count(void);
void fluid_launch_count_reset(void);

/* Modes used by the Stam-signature host-pointer entry points below (their
 * argument lists have no room for them).  Defaults: FLUID_GS_WAVEFRONT, 20,
 * FLUID_ADVECT_MIDPOINT -- the reference's behaviour.  Thread-local. */
int fluid_set_gs_order(int order);
int fluid_set_gs_iters(int iters);
int fluid_set_advect_mode(int mode);
/* When 0, vel_step/dens_step/project/diffuse_blur do not copy the arrays the
 * reference only clobbers as scratch (u0,v0,x0,p,div) back to the host.
 * Default 1 = every array the reference writes is written. */
int fluid_set_scratch_download(int on);

/* CUDA device used by the host-pointer entry points and by contexts created
 * with device = -1 on the calling thread (default: device 0). */
int fluid_set_device(int device);

/* Pinned host memory for callers that want full-speed transfers. */
int fluid_host_alloc(void **p, size_t bytes);
int fluid_host_free(void *p);

/* ---- host-pointer entry points: the reference's primitives ------------- *
 * Synchronous; arrays are dense (N+2)^2 host floats; copy in, run on the    *
 * GPU, copy out; nothing is retained.  `list` (std::vector<float>&,         *
 * FluidCode.h:4) becomes float*; raw-pointer twins of two of them already   *
 * exist in the reference (main.cpp:24,43).                                  */

/* AddArrDt, FluidSystem.cpp:449-453: x += dt*s on all (N+2)^2 cells. */
int fluid_add_source(int N, float *x, const float *s, float dt);
/* SetBorder, FluidSystem.cpp:465-500. */
int fluid_set_bnd(int N, int b, float *x);
/* LinearSolve1, FluidSystem.cpp:534-548: a>0 ? gs_iters : 1 sweeps, SetBorder each. */
int fluid_lin_solve(int N, int b, float *x, const float *x0, float a, float c);
/* Diffuse1, FluidSystem.cpp:810-814. */
int fluid_diffuse(int N, int b, float *x, const float *x0, float diff, float dt);
/* Diffuse_blur -> Blur, FluidSystem.cpp:833-837, 722-743: result in x, the
 * penultimate iterate in x0 (x0 is clobbered, as in the reference). */
int fluid_diffuse_blur(int N, int b, float *x, float *x0, float diff, float dt);
/* Advect1, FluidSystem.cpp:850-862; d0 may alias u or v. */
int fluid_advect(int N, int b, float *d, const float *d0, const float *u, const float *v, float dt);
/* Project, FluidSystem.cpp:982-1007; p and div are scratch outputs. */
int fluid_project(int N, float *u, float *v, float *p, float *div);
/* VelStep1, FluidSystem.cpp:354-376; u0,v0 = sources in, scratch out. */
int fluid_vel_step(int N, float *u, float *v, float *u0, float *v0, float visc, float dt);
/* DensStep1 (one fluid), FluidSystem.cpp:289-299; x0 = source in, scratch out. */
int fluid_dens_step(int N, float *x, float *x0, const float *u, const float *v, float diff, float dt);
/* FluidSystem::Update, FluidSystem.cpp:101-128 -- the reference's public per-frame call
 * (FSType::ORIGINAL, one fluid): zero the source arrays, apply the forcing list (what the
 * updaters do, FluidSystem.cpp:118,278-287), [vorticity], vel step, dens step, [damp].
 * Only u, v and dens cross the bus (3 arrays up, 3 down). */
int fluid_update(int N, float *u, float *v, float *dens, int nsrc, const fluid_source *src, float visc, float diff,
		 float vorticity, float damp, float dt);
/* VortConfinement, FluidSystem.cpp:1124-1175. */
int fluid_vort_confinement(int N, float *u, float *v, float vorticity);
/* Damp, FluidSystem.cpp:1177-1183: vel *= powf(1-amount, dt), factor computed on the host. */
int fluid_damp(int N, float *vel, float amount, float dt);

/* ---- resident context: the timed path ---------------------------------- *
 * Plays the role of a FluidSystem set up with FSType::ORIGINAL and one      *
 * fluid (FluidSystem.cpp:18-99): owns the six fields on the device.         */

void fluid_params_default(fluid_params *p);
int fluid_ctx_create(const fluid_params *p, fluid_ctx **out);
int fluid_ctx_destroy(fluid_ctx *ctx);
/* FluidSystem::Clear, FluidSystem.cpp:163-170. */
int fluid_ctx_clear(fluid_ctx *ctx);
/* Dense host array <-> device field; `sim` selects the ensemble member. */
int fluid_ctx_upload(fluid_ctx *ctx, int field, int sim, const float *host);
int fluid_ctx_download(fluid_ctx *ctx, int field, int sim, float *host);
/* Forcing applied at the start of every step; n entries for ensemble member
 * `sim` (sim = -1: the same list for every member). */
int fluid_ctx_set_sources(fluid_ctx *ctx, int sim, int n, const fluid_source *src);
/* pVelStep / pDensStep, FluidSystem.cpp:120,124 (no zero fill, no sources). */
int fluid_ctx_vel_step(fluid_ctx *ctx, float dt);
int fluid_ctx_dens_step(fluid_ctx *ctx, float dt);
/* FluidSystem::Update, FluidSystem.cpp:101-128, nsteps times: zero the Prev
 * fields, apply sources, [vorticity], vel step, dens step, [damp].  Asynchronous. */
int fluid_ctx_step(fluid_ctx *ctx, float dt, int nsteps);
/* Block until everything queued on the context has finished. */
int fluid_ctx_sync(fluid_ctx *ctx);
/* Device time of the work queued by the last fluid_ctx_step call, in ms
 * (CUDA events on the context's stream; waits for completion). */
int fluid_ctx_last_step_ms(fluid_ctx *ctx, float *ms);
/* Change modes / coefficients of a live context (steps, visc... are public
 * members of FluidSystem, FluidSystem.h:142-152). */
int fluid_ctx_set_modes(fluid_ctx *ctx, int gs_order, int gs_iters, int advect_mode);
int fluid_ctx_set_coeffs(fluid_ctx *ctx, float visc, float diff, float vorticity, float damp);
/* FluidProps of fluid k (FluidCode.h:46-59): its diffusion coefficient, and its viscosity, which
 * enters the velocity step through visc_avg = mean over the fluids (FluidSystem.cpp:26-30).
 * fluid_ctx_set_coeffs / fluid_params set fluid 0 and are copied to the others until this is called. */
int fluid_ctx_set_fluid(fluid_ctx *ctx, int k, float diff, float visc);
/* Kernel selection for FLUID_GS_REDBLACK: sweeps fused per shared-memory tile
 * residency (2 = default, 1; 0 = one kernel launch per colour).  Results are
 * bit-identical for every setting; only the memory traffic differs. */
int fluid_ctx_set_tuning(fluid_ctx *ctx, int rb_tile_sweeps);
/* Large grids: the register-resident streaming kernels -- red-black relaxation, 4 sweeps per launch, and Blur,
 * 5 passes per launch -- with the shared-memory tile kernel / the pass-per-launch kernel as their exact fallback
 * (1, default; 0 = tile and pass kernels everywhere).  Dividends outside the proven window of the short division
 * (the subnormal fringe of a field) are normally listed per segment and recomputed exactly behind the launch; when
 * most segments come back (ensembles of small grids) the library switches by itself to the scaled division
 * sequence, which is exact on the fringe as well (proven per divisor like the window; nine instructions more per
 * update); 2 forces that sequence.  Results identical. */
int fluid_ctx_set_stream_kernel(fluid_ctx *ctx, int on);
/* fluid_ctx_step / fluid_update / fluid_ctx_vel_step: run independent parts side by side on extra streams -- the
 * densities' add_source + Blur beside VelStep, and VelStep's v diffusion beside its u diffusion (which costs one
 * more scratch field) -- (default on; 0 = one stream, kernels strictly in the
 * reference's order -- what per-kernel profiling wants).  Results identical. */
int fluid_ctx_set_overlap(fluid_ctx *ctx, int on);
/* Advect1 (FluidSystem.cpp:850-862) kernel choice: 0 = automatic (row-window kernel on grids of 2^20 cells or
 * more), 1 = one thread per cell with global gathers, 2 = row window (a warp keeps rows j-1..j+1 of the sampled
 * fields in registers; backtraces shorter than one cell interpolate from registers, longer ones gather).
 * Results identical. */
int fluid_ctx_set_advect_kernel(fluid_ctx *ctx, int which);
/* fluid_ctx_step / fluid_update: the source arrays of a frame are zero except for the listed PREV_SET cells
 * (Update zero-fills them, FluidSystem.cpp:104-109), so AddArrDt (:449-453) over them is known in advance: every
 * value stays, a -0 becomes +0, the listed cells get x + dt*source.  1 (default) = do exactly that (one read of
 * the field, almost no writes) on grids larger than 128^2; 0 = zero-fill and add densely, as the reference does.
 * Results identical. */
int fluid_ctx_set_sparse_sources(fluid_ctx *ctx, int on);
/* ---- cell map of a FLUID_TYPE_ORIGINAL_BORDERED context ------------------ *
 * The reference rebuilds cellInfo at the top of every Update: ResetCellInfo, colliders, CalcBorderFromContent,
 * borders (FluidSystem.cpp:112-115).  Without moving objects that yields the same map every frame, so a context
 * keeps ONE map and uses it until the host changes it.  A new context holds ResetCellInfo + CalcBorderFromContent
 * (all interior cells FLUID, walls along the outer ring).  Dense host maps are (N+2)^2 bytes, ID(i,j) order. */
/* ResetCellInfo + CalcBorderFromContent (FluidSystem.cpp:1215-1232, 1185-1213). */
int fluid_ctx_reset_cell_info(fluid_ctx *ctx);
/* DisableEdge (FluidSystem.cpp:1388-1408) for n edges, 4 ints each: x0, y0, x1, y1 (adjacent cells). */
int fluid_ctx_disable_edges(fluid_ctx *ctx, int n, const int *edges);
/* SquareBorder::WriteBorder (FluidBorder.cpp:4-19): the shipped scene's closed square wall around the centre. */
int fluid_ctx_add_square_border(fluid_ctx *ctx);
/* Replace / read the whole map (what a collider or a custom FluidBorder would have produced, FluidSystem.h GetInfo).
 * recalc != 0: run CalcBorderFromContent on the uploaded FLUID / OBJECT bits afterwards. */
int fluid_ctx_set_cell_info(fluid_ctx *ctx, const unsigned char *host_info, int recalc);
int fluid_ctx_get_cell_info(fluid_ctx *ctx, unsigned char *host_info);
/* FLUID_TYPE_ORIGINAL_BORDERED_MF: viscosity of cells that hold no fluid (visc_air, FluidSystem.h:149; default 0.001). */
int fluid_ctx_set_visc_air(fluid_ctx *ctx, float visc_air);
/* Tile edge of the reference-order (wavefront) kernel: 0 = automatic, 32 or 64.  Results identical. */
int fluid_ctx_set_wavefront_tile(fluid_ctx *ctx, int tile);
/* Reference-order solve as ONE launch: 0 = never (one launch per hyperplane of the (tile, sweep) schedule),
 * 2 = CTAs take (sweep, tile column) tickets and order themselves with progress counters, 3 = the row pipeline (a
 * warp per band of 30 rows and sweep streams along the rows; bands and sweeps hand over through counters),
 * 1 = automatic (default): whichever measured fastest for the grid's size.  Results identical. */
int fluid_ctx_set_wavefront_persistent(fluid_ctx *ctx, int on);
/* Grids of N <= 128 run a whole lin_solve (all sweeps, either order) or a whole Blur
 * (all passes) as ONE single-CTA launch per ensemble member, from shared memory.  on = 0
 * forces the large-grid kernels (identical results; used by the tests). */
int fluid_ctx_set_small_grid_kernels(fluid_ctx *ctx, int on);
/* A solve divides by the same c throughout.  For a given c the library PROVES on the GPU over which range of
 * dividends a 3-instruction multiply/fma sequence equals the IEEE quotient bit for bit, and uses it there (exact
 * division elsewhere).  mode 0 = never; 1 (default) = the quick proof at a divisor's first use -- the sequence
 * commutes with power-of-two scaling of the dividend, so one binade of dividends decides it: ~0.1 ms and one event
 * wait per new divisor, cached per context; 2 = the exhaustive proof (all 2^31 non-negative finite dividends, ~3 ms),
 * at once and waited for (tests).  Results never change. */
int fluid_ctx_set_fast_division(fluid_ctx *ctx, int mode);
/* The proven window for `divisor`: dividends with 2^lo_exp <= |t| < 2^(hi_exp+1)
 * (hi_exp < lo_exp: none).  Runs the proof if it has not been run yet. */
int fluid_ctx_division_window(fluid_ctx *ctx, float divisor, int *lo_exp, int *hi_exp);

/* ---- row slabs over several GPUs ---------------------------------------- *
 * One context per GPU/process, each a slab (fluid_params.slab_*).  No          *
 * counterpart in the reference (single process, single thread).  Slabs overlap *
 * by slab_halo+1 rows and recompute the overlap, so halos are refreshed once    *
 * per 20-sweep solve (6 times per step) rather than once per colour pass.  The  *
 * library decides WHEN; the host supplies HOW: the callback must, for each      *
 * listed field, move fluid_ctx_halo_regions()' send blocks into the             *
 * neighbours' recv blocks (ncclSend/ncclRecv, cudaMemcpyPeer ...) and return 0   *
 * only when this rank's recv blocks are complete.  The context's stream is      *
 * idle when the callback runs.  upload/download take the GLOBAL dense array.    *
 * Forcing entries use global rows.  FLUID_GS_REDBLACK only.                     */
typedef int (*fluid_exchange_fn)(void *user, int nfields, const int *fields);
/* One halo refresh of 1 or 2 fields by themselves, through whichever transport the slab has (what the steps do at their
 * six refresh points; for tests and for timing a transport).  Every slab must make the same call. */
int fluid_ctx_halo_refresh(fluid_ctx *ctx, int nfields, const int *fields);
int fluid_ctx_set_exchange(fluid_ctx *ctx, fluid_exchange_fn fn, void *user);
/* Native transport: halo rows pushed straight into the neighbour's memory over NVLink and ordered by flag words,
 * two small kernels per refresh on the context's stream (csrc/fluid_peer.cu) -- no callback, no host round trip,
 * no collective.  Each slab owns one block (inboxes + flags):
 *   fluid_ctx_peer_halo_local    allocates it (once) and returns its device address and size;
 *   fluid_ctx_peer_halo_connect  takes the addresses of the lower / upper neighbour's blocks AS THIS PROCESS SEES
 *                                THEM (NULL at the two ends): directly when all slabs live in one process with peer
 *                                access enabled, else a mapping made with fluid_ipc_import from the handle the
 *                                neighbour's process obtained with fluid_ipc_export (CUDA IPC, 64 bytes, to be sent
 *                                over whatever channel the host has -- MPI, torch.distributed, a pipe).
 * Every slab must be connected before any of them steps, all slabs must run the same sequence of calls, and the
 * slabs must sit on DIFFERENT GPUs (a slab's pull kernel waits for its neighbours' push kernels).  A block must
 * outlive every neighbour's last step. */
int fluid_ctx_peer_halo_local(fluid_ctx *ctx, void **block, size_t *bytes);
int fluid_ctx_peer_halo_connect(fluid_ctx *ctx, void *lower_block, void *upper_block);
int fluid_ipc_export(void *device_ptr, unsigned char handle[64]);
int fluid_ipc_import(const unsigned char handle[64], void **device_ptr);
int fluid_ipc_close(void *device_ptr);
/* Run the context on the caller's stream (a cudaStream_t; NULL = back to the
 * context's own).  With exchange_stream_ordered != 0 the library does not wait for
 * the stream before calling the exchange callback: the transport must then enqueue
 * its transfers in stream order (e.g. NCCL on that stream), and the whole step stays
 * asynchronous -- no host round trip per halo exchange. */
int fluid_ctx_set_stream(fluid_ctx *ctx, void *stream, int exchange_stream_ordered);
/* Device pointers of the blocks to move for `field`; *count floats each; a
 * pointer is NULL where there is no neighbour.  send_down -> lower neighbour's
 * recv_up, send_up -> upper neighbour's recv_down.  Ask at every exchange: a
 * red-black solve may leave a field in another device array than before (it
 * trades arrays with the solver's scratch field instead of copying), so pointers
 * obtained earlier can be stale. */
int fluid_ctx_halo_regions(fluid_ctx *ctx, int field, float **send_down, float **recv_down, float **send_up, float **recv_up,
			   size_t *count);

/* Several GPUs from one plain-C process: `ndev` row slabs of p->slab_rows rows each (global grid
 * N x slab_rows*ndev) on devices[0..ndev), stepped together; halos move by peer copies
 * (cudaMemcpyPeerAsync over NVLink) ordered by CUDA events only -- no stream is drained and the
 * host never waits for the device.  A device may be listed more than once (used by the tests to
 * run the multi-GPU path on one GPU).  upload/download take the GLOBAL dense array, forcing
 * entries use global rows, relaxation is FLUID_GS_REDBLACK.  No counterpart in the reference. */
typedef struct fluid_multi fluid_multi;
int fluid_multi_create(const fluid_params *p, int ndev, const int *devices, fluid_multi **out);
int fluid_multi_destroy(fluid_multi *m);
int fluid_multi_count(fluid_multi *m);
int fluid_multi_set_sources(fluid_multi *m, int n, const fluid_source *src);
int fluid_multi_upload(fluid_multi *m, int field, const float *host_global);
int fluid_multi_download(fluid_multi *m, int field, float *host_global);
int fluid_multi_step(fluid_multi *m, float dt, int nsteps);   /* returns when the work is enqueued */
int fluid_multi_sync(fluid_multi *m);
int fluid_multi_last_step_ms(fluid_multi *m, float *ms);     /* slowest device */

/* Slab-local transfers: `host` holds exactly the rows this context stores (halo
 * rows included), dense, (N+2) floats per row; fluid_ctx_local_rows tells how
 * many rows that is and which global row the first one is. */
int fluid_ctx_upload_local(fluid_ctx *ctx, int field, const float *host);
int fluid_ctx_download_local(fluid_ctx *ctx, int field, float *host);
int fluid_ctx_local_rows(fluid_ctx *ctx, int *rows_stored, int *first_global_row);

/* Headless display path: the colourise loop and the velocity line list of
 * FluidRenderer::DoBuffer (FluidRenderer.cpp:56-84, 97-132) computed on the device and copied to dense host
 * arrays: host_pixels (N+2)^2 x 3 floats, host_lines (N+2)^2 x 4 floats or NULL.  color_rgb holds 3 floats per
 * fluid of the context (FluidProps::color, FluidCode.h:50): FLUID cells get every fluid's colour weighted by its
 * density, normalised where the summed density exceeds 1; OBJECT cells (bordered contexts' cell map) are
 * magenta, all other cells green.  No window, texture upload or GL interop: graphics APIs are unavailable on
 * this box. */
int fluid_ctx_render(fluid_ctx *ctx, int sim, const float *color_rgb, float *host_pixels, float *host_lines);

/* Per-primitive device time inside fluid_ctx_step / vel_step / dens_step: when
 * on, every primitive is bracketed by a CUDA event pair on the context's
 * stream.  fluid_ctx_profile waits for the stream, then sums milliseconds and
 * kernel launches per class (arrays of FLUID_PROF_COUNT) since the last reset.
 * (The reference's only timing is the frame-level TimeAnalyse,
 * SystemManager.cpp:246-275.) */
typedef enum fluid_prof_class {
	FLUID_PROF_LIN_SOLVE = 0,   /* Gauss-Seidel sweeps (wavefront or red-black) */
	FLUID_PROF_BLUR = 1,        /* Jacobi blur passes incl. their border launches */
	FLUID_PROF_ADVECT = 2,
	FLUID_PROF_PROJECT_AUX = 3, /* divergence + gradient subtraction */
	FLUID_PROF_ADD_SOURCE = 4,
	FLUID_PROF_OTHER = 5,       /* zero fill, forcing, vorticity, damp */
	FLUID_PROF_EXCHANGE = 6,    /* halo exchange between row slabs (includes the stream sync before it) */
	FLUID_PROF_COUNT = 7
} fluid_prof_class;
int fluid_ctx_set_profiling(fluid_ctx *ctx, int on);
int fluid_ctx_profile(fluid_ctx *ctx, float *ms, unsigned long long *launches, int reset);

/* Single primitives on resident fields (x, x0 ... are fluid_field ids; every
 * ensemble member is processed).  Used by the parity tests and by bench.py's
 * per-kernel roofline measurement. */
int fluid_ctx_add_source(fluid_ctx *ctx, int x, int s, float dt);
int fluid_ctx_set_bnd(fluid_ctx *ctx, int b, int x);
int fluid_ctx_lin_solve(fluid_ctx *ctx, int b, int x, int x0, float a, float c);
int fluid_ctx_blur(fluid_ctx *ctx, int b, int x, int x0, float a, float c);
int fluid_ctx_advect(fluid_ctx *ctx, int b, int d, int d0, int u, int v, float dt);
int fluid_ctx_advect_velocity(fluid_ctx *ctx, int u, int v, int u0, int v0, float dt);
int fluid_ctx_project(fluid_ctx *ctx, int u, int v, int p, int div);
/* Time `reps` back-to-back repetitions of one primitive with CUDA events on
 * the context's stream; *ms = mean per repetition.  which: see fluid_kernel_id. */
typedef enum fluid_kernel_id {
	FLUID_K_LIN_SOLVE = 0, FLUID_K_BLUR = 1, FLUID_K_ADVECT = 2, FLUID_K_ADVECT_VEL = 3,
	FLUID_K_PROJECT = 4, FLUID_K_ADD_SOURCE = 5
} fluid_kernel_id;
int fluid_ctx_time_kernel(fluid_ctx *ctx, int which, int reps, float *ms);

/* ---- for callers that must not allocate --------------------------------- *
 * No counterpart in the reference (it owns all its memory as std::vector and *
 * never allocates inside a step either).                                     *
 *                                                                            *
 * Red-black lin_solve (same sweeps, same results as fluid_lin_solve with     *
 * FLUID_GS_REDBLACK) on DEVICE arrays you already own, without a single      *
 * cudaMalloc/cudaFree: meant for real-time loops, plug-ins living inside a   *
 * pre-sized arena, or processes whose GPU memory belongs to a pool allocator *
 * -- places where a hidden allocation can fail late, stall the device, or is *
 * simply forbidden.  The tile ping-pong scratch that a context would         *
 * allocate is a fixed array compiled into the library (4.3 MiB resident on   *
 * every GPU that loads it), hence the limit N <= FLUID_STATIC_MAX_N; larger  *
 * N is refused with FLUID_ERR_INVALID.                                       *
 *                                                                            *
 * Layout of x_dev and x0_dev (16-byte aligned, on the current CUDA device):  *
 * (N+2) rows of FLUID_STATIC_PITCH_OF(N) floats, cell (i,j) at               *
 * j*pitch + FLUID_STATIC_XO + i; the other floats of a row are padding.      *
 * x_dev is updated in place; x0_dev is only read.  `stream` is a             *
 * cudaStream_t (NULL = default stream); the call only enqueues work.         *
 *                                                                            *
 * - The first call on each device looks up the array's device address and    *
 *   raises the kernels' shared-memory limit; make that first call before any *
 *   stream capture begins, and from ONE host thread: the per-device set-up   *
 *   record is written without a lock, so two first calls racing on the same  *
 *   device from different threads are not supported.                         *
 * - All calls share the one array, and it keeps the previous call's data     *
 *   (this function clears the part it is about to use).  Two calls whose GPU *
 *   work can overlap -- different streams, different host threads -- race on *
 *   it; keep them on one stream or serialise them yourself.                  */
#define FLUID_STATIC_MAX_N 1024
#define FLUID_STATIC_XO 31
#define FLUID_STATIC_PITCH_OF(N) ((((N) + 2 + FLUID_STATIC_XO + 31) / 32) * 32)
int fluid_lin_solve_rb_static(int N, int b, float *x_dev, const float *x0_dev, float a, float c, int iters, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* FLUID_CUDA_H */
