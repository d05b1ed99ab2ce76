// TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called from the
// product (libfluid_cuda.so / fluid_b200).  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load the library this
// file builds (oracle/_ref/libfluid_ref.so).
//
// A flat C API over the UNMODIFIED reference solver: this translation unit and
// the reference's own FluidSystem.cpp / FluidUpdaters.cpp / FluidBorder.cpp are
// compiled where they lie under /root/reference (see oracle/Makefile); no
// reference source is copied into this repository.  All arithmetic executed
// through these entry points is the reference's.
//
// Private solver primitives (FluidSystem.h:71-140) are reached with
// `#define private public` placed AFTER every standard header.
#include <Tools/Include.h>
#include <Tools/EigenInc.h>
#include <Tools/Math.h>
#include <cstring>
#include <future>
#include <limits>
#include <sstream>

#define private public
#define protected public
#include "FluidSystem.h"
#include "FluidCollider.h"
#include "FluidBorder.h"
#undef private
#undef protected

// FluidCollider.cpp needs real Eigen and cannot be compiled here; the ORIGINAL
// pipeline never registers a collider, so these are never reached
// (FluidSystem.cpp:1234-1254 iterate an empty gColliders).
FluidCollider::FluidCollider() {}
FluidCollider::~FluidCollider() {}
void FluidCollider::Update(int, FluidSystem*, std::vector<byte>&, real) {}
void FluidCollider::ApplyForces(int, real, FluidSystem*) {}
void FluidCollider::Collide(FluidCollider*, FluidCollider*) {}

namespace {

struct SourceEntry { int field, i, j, mode; float value; };

// Synthetic forcing, injected the way the reference's own updaters are
// (Tools::UpdatableR<bool, FluidUpdate>, FluidSystem.cpp:278-287).
// field: 0=u 1=v 2+k=density of fluid k;  mode: 0 = write Prev (Stam-style source, relies on
// Update's zero fill, FluidSystem.cpp:104-109), 1 = Curr +=, 2 = Curr =
// (the last two are what Blower/AddFluid do, FluidSystem.cpp:250-257).
class SourceUpdater : public Tools::UpdatableR<bool, FluidUpdate> {
public:
	std::vector<SourceEntry> entries;
	bool Update(FluidUpdate fu) override
	{
		if (fu.type != FluidUpdate::VEL) return true;
		FluidSystem* fs = fu.fs;
		for (const SourceEntry& e : entries) {
			FlipFlopArr<real>& f = e.field == 0 ? fs->vX : (e.field == 1 ? fs->vY : fs->densList[e.field - 2]);
			int id = fs->ID(e.i, e.j);
			if (e.mode == 0) f.Prev()[id] = e.value;
			else if (e.mode == 1) f.Curr()[id] += e.value;
			else f.Curr()[id] = e.value;
		}
		return true;
	}
};

struct RefSim {
	FluidSystem fs;
	SourceUpdater* src = nullptr; // owned by fs.gUpdaters semantics; we delete it ourselves
};

FluidSystem* scratch_system(int N, int steps)
{
	// A FluidSystem whose only role is to carry N and `steps` for a primitive call.
	FluidSystem* fs = new FluidSystem();
	fs->N = N;
	fs->steps = steps;
	fs->fluid_count = 0;
	return fs;
}

typedef std::vector<real> vec;
inline vec load(const float* p, int N) { return vec(p, p + (size_t)(N + 2) * (N + 2)); }
inline void store(float* p, const vec& v) { std::memcpy(p, v.data(), v.size() * sizeof(float)); }

} // namespace

extern "C" {

void* ref_create(int N, float visc, float diff, float vorticity, int steps, int use_vort, float damp)
{
	RefSim* s = new RefSim();
	std::vector<FluidProps> props = { { { diff, visc }, { 1.0f, 1.0f, 1.0f } } };
	s->fs.Setup(N, vorticity, props, FluidSystem::FSType::ORIGINAL);
	s->fs.steps = steps;
	s->fs.useVort = use_vort != 0;
	s->fs.damp = damp;
	s->fs.Clear();
	std::fill(s->fs.p.begin(), s->fs.p.end(), 0.0f);
	s->src = new SourceUpdater();
	s->fs.AddUpdater(s->src);
	return s;
}

// Several fluids, each with its own FluidProps (FluidCode.h:46-59): what the shipped scene sets up
// (SystemManager.cpp:103-108), but on the ORIGINAL pipeline.
void* ref_create_multi(int N, int nfluids, const float* diffs, const float* viscs, float vorticity, int steps, int use_vort,
		       float damp)
{
	RefSim* s = new RefSim();
	std::vector<FluidProps> props;
	for (int k = 0; k < nfluids; k++) props.push_back({ { diffs[k], viscs[k] }, { 1.0f, 1.0f, 1.0f } });
	s->fs.Setup(N, vorticity, props, FluidSystem::FSType::ORIGINAL);
	s->fs.steps = steps;
	s->fs.useVort = use_vort != 0;
	s->fs.damp = damp;
	s->fs.Clear();
	std::fill(s->fs.p.begin(), s->fs.p.end(), 0.0f);
	s->src = new SourceUpdater();
	s->fs.AddUpdater(s->src);
	return s;
}

// Any solver type (FluidSystem.h:10-17: 0 ORIGINAL, 1 ORIGINAL_BORDERED, 2 ORIGINAL_BORDERED_MF) with
// per-fluid properties and, optionally, the shipped scene's internal wall (SquareBorder, FluidBorder.cpp:4-19).
void* ref_create_typed(int N, int nfluids, const float* diffs, const float* viscs, float vorticity, int steps, int use_vort,
		       float damp, int fstype, int square_border)
{
	RefSim* s = new RefSim();
	std::vector<FluidProps> props;
	for (int k = 0; k < nfluids; k++) props.push_back({ { diffs[k], viscs[k] }, { 1.0f, 1.0f, 1.0f } });
	s->fs.Setup(N, vorticity, props, (FluidSystem::FSType) fstype);
	s->fs.steps = steps;
	s->fs.useVort = use_vort != 0;
	s->fs.damp = damp;
	s->fs.Clear();
	std::fill(s->fs.p.begin(), s->fs.p.end(), 0.0f);
	s->src = new SourceUpdater();
	s->fs.AddUpdater(s->src);
	if (square_border) s->fs.AddBorder(new SquareBorder()); // (leaked with the simulation: FluidSystem has no border removal)
	return s;
}

// the system's own pressure array (FluidSystem.h:162; ProjectBO solves into it, FluidSystem.cpp:1096-1098)
float* ref_p(void* h) { return ((RefSim*) h)->fs.p.data(); }
float ref_visc_air(void* h) { return ((RefSim*) h)->fs.visc_air; }

// cellInfo as the last Update (or Setup) left it (FluidSystem.h: GetInfo)
unsigned char* ref_cell_info(void* h) { return ((RefSim*) h)->fs.cellInfo.data(); }

float* ref_density(void* h, int fluid, int prev)
{
	RefSim* s = (RefSim*) h;
	return prev ? s->fs.densList[fluid].Prev().data() : s->fs.densList[fluid].Curr().data();
}

void ref_destroy(void* h)
{
	RefSim* s = (RefSim*) h;
	s->fs.gUpdaters.clear();
	delete s->src;
	delete s;
}

void ref_set_sources(void* h, int n, const int* field, const int* i, const int* j, const int* mode, const float* value)
{
	RefSim* s = (RefSim*) h;
	s->src->entries.clear();
	for (int k = 0; k < n; k++) s->src->entries.push_back({ field[k], i[k], j[k], mode[k], value[k] });
}

// which: 0=u.Curr 1=v.Curr 2=dens.Curr 3=u.Prev 4=v.Prev 5=dens.Prev
float* ref_field(void* h, int which)
{
	RefSim* s = (RefSim*) h;
	switch (which) {
	case 0: return s->fs.vX.Curr().data();
	case 1: return s->fs.vY.Curr().data();
	case 2: return s->fs.densList[0].Curr().data();
	case 3: return s->fs.vX.Prev().data();
	case 4: return s->fs.vY.Prev().data();
	case 5: return s->fs.densList[0].Prev().data();
	}
	return nullptr;
}

// FluidSystem::Update, FluidSystem.cpp:101-128 (the whole frame).
void ref_update(void* h, float dt, int nsteps)
{
	RefSim* s = (RefSim*) h;
	for (int k = 0; k < nsteps; k++) s->fs.Update(dt);
}

// VelStep1 / DensStep1 alone (FluidSystem.cpp:354-376, 289-299), no zero fill.
void ref_vel_step(void* h, float dt) { ((RefSim*) h)->fs.VelStep1(dt); }
void ref_dens_step(void* h, float dt) { ((RefSim*) h)->fs.DensStep1(dt); }
void ref_vort_confinement(void* h) { ((RefSim*) h)->fs.VortConfinement(); }
void ref_damp(void* h, float dt)
{
	RefSim* s = (RefSim*) h;
	s->fs.Damp(s->fs.vX.Curr(), s->fs.damp, dt);
	s->fs.Damp(s->fs.vY.Curr(), s->fs.damp, dt);
}
float ref_total_inner_mass(void* h) { return ((RefSim*) h)->fs.TotalInnerMass(true); }

// ---- primitives on caller buffers (dense (N+2)^2, ID = i + (N+2) j) -------

void ref_add_source(int N, float* x, const float* s, float dt)
{
	FluidSystem* fs = scratch_system(N, 20);
	vec vx = load(x, N), vs = load(s, N);
	fs->AddArrDt(N, vx, vs, dt);
	store(x, vx);
	delete fs;
}

void ref_set_bnd(int N, int b, float* x)
{
	FluidSystem* fs = scratch_system(N, 20);
	vec vx = load(x, N);
	fs->SetBorder(N, b, vx);
	store(x, vx);
	delete fs;
}

void ref_lin_solve(int N, int b, float* x, const float* x0, float a, float c, int steps)
{
	FluidSystem* fs = scratch_system(N, steps);
	vec vx = load(x, N), v0 = load(x0, N);
	fs->LinearSolve1(N, b, vx, v0, a, c);
	store(x, vx);
	delete fs;
}

void ref_diffuse(int N, int b, float* x, const float* x0, float diff, float dt, int steps)
{
	FluidSystem* fs = scratch_system(N, steps);
	vec vx = load(x, N), v0 = load(x0, N);
	fs->Diffuse1(N, b, vx, v0, diff, dt);
	store(x, vx);
	delete fs;
}

// Blur swaps the two vectors' buffers an even number of times (FluidSystem.cpp:730),
// so the result is in x and the penultimate iterate in x0: both are copied back.
void ref_diffuse_blur(int N, int b, float* x, float* x0, float diff, float dt, int steps)
{
	FluidSystem* fs = scratch_system(N, steps);
	vec vx = load(x, N), v0 = load(x0, N);
	fs->Diffuse_blur(N, b, vx, v0, diff, dt);
	store(x, vx);
	store(x0, v0);
	delete fs;
}

void ref_advect(int N, int b, float* d, const float* d0, const float* u, const float* v, float dt)
{
	FluidSystem* fs = scratch_system(N, 20);
	vec vd = load(d, N), vd0 = load(d0, N), vu = load(u, N), vv = load(v, N);
	fs->Advect1(N, b, vd, vd0, vu, vv, dt);
	store(d, vd);
	delete fs;
}

// Stam's advect (single Euler backtrace): the reference keeps it as the
// uncalled EulerBack (FluidSystem.h:208-219); loop + SetBorder as Advect1.
void ref_advect_euler(int N, int b, float* d, const float* d0, const float* u, const float* v, float dt)
{
	FluidSystem* fs = scratch_system(N, 20);
	vec vd = load(d, N), vd0 = load(d0, N), vu = load(u, N), vv = load(v, N);
	real dtN = dt * N;
	for (int j = 1; j <= N; j++)
		for (int i = 1; i <= N; i++)
			fs->EulerBack(N, i, j, dtN, vu, vv, vd, vd0);
	fs->SetBorder(N, b, vd);
	store(d, vd);
	delete fs;
}

// ---- bordered variants (FluidSystem.cpp:550-578, 745-777, 864-874, 1009-1059, 1185-1232, 1388-1408) on caller buffers ----

static FluidSystem* scratch_system_info(int N, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system(N, steps);
	fs->cellInfo.assign(info, info + (size_t) (N + 2) * (N + 2));
	return fs;
}

// ResetCellInfo + CalcBorderFromContent [+ SquareBorder::WriteBorder], i.e. what Update does to cellInfo
// each frame when no collider is registered (FluidSystem.cpp:112-115).
void ref_make_cell_info(int N, int square_border, unsigned char* info)
{
	FluidSystem* fs = scratch_system(N, 20);
	fs->cellInfo.resize((size_t) (N + 2) * (N + 2));
	fs->ResetCellInfo();
	fs->CalcBorderFromContent();
	if (square_border) {
		SquareBorder sb;
		sb.WriteBorder(N, fs, fs->cellInfo);
	}
	std::memcpy(info, fs->cellInfo.data(), fs->cellInfo.size());
	delete fs;
}

// CalcBorderFromContent alone on a caller-made map (FLUID / OBJECT bits chosen by the caller)
void ref_calc_border(int N, unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, 20, info);
	fs->CalcBorderFromContent();
	std::memcpy(info, fs->cellInfo.data(), fs->cellInfo.size());
	delete fs;
}

void ref_lin_solve_b(int N, int b, float* x, const float* x0, float a, float c, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, steps, info);
	vec vx = load(x, N), v0 = load(x0, N);
	fs->LinearSolve1B(N, b, vx, v0, a, c);
	store(x, vx);
	delete fs;
}

void ref_diffuse_blur_b(int N, int b, float* x, float* x0, float diff, float dt, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, steps, info);
	vec vx = load(x, N), v0 = load(x0, N);
	fs->Diffuse_blurB(N, b, vx, v0, diff, dt);
	store(x, vx);
	store(x0, v0);
	delete fs;
}

void ref_advect_b(int N, int b, float* d, const float* d0, const float* u, const float* v, float dt)
{
	FluidSystem* fs = scratch_system(N, 20);
	vec vd = load(d, N), vd0 = load(d0, N), vu = load(u, N), vv = load(v, N);
	fs->Advect1B(N, b, vd, vd0, vu, vv, dt);
	store(d, vd);
	delete fs;
}

void ref_project_b(int N, float* u, float* v, float* p, float* div, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, steps, info);
	vec vu = load(u, N), vv = load(v, N), vp = load(p, N), vd = load(div, N);
	fs->ProjectB(N, vu, vv, vp, vd);
	store(u, vu); store(v, vv); store(p, vp); store(div, vd);
	delete fs;
}

// LinearSolve1BO / ProjectBO (FluidSystem.cpp:580-615, 1061-1122) on caller buffers; `p` is loaded into and
// read back from the member array the reference really uses.
void ref_lin_solve_bo(int N, int b, float* x, const float* x0, float a, float c, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, steps, info);
	vec vx = load(x, N), v0 = load(x0, N);
	fs->LinearSolve1BO(N, b, vx, v0, a, c);
	store(x, vx);
	delete fs;
}

void ref_project_bo(int N, float* u, float* v, float* p, float* div, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, steps, info);
	vec vu = load(u, N), vv = load(v, N), vd = load(div, N), unused = load(p, N);
	fs->p = load(p, N);
	fs->ProjectBO(N, vu, vv, unused, vd);
	store(u, vu); store(v, vv); store(p, fs->p); store(div, vd);
	delete fs;
}

// LinearSolve1BSO (FluidSystem.cpp:673-720) with nfluids density fields and their coefficients of one kind
void ref_lin_solve_bso(int N, int b, float* x, const float* x0, float aSc, float a_fallback, int nfluids, const float* const* dens,
		       const float* coef, int steps, const unsigned char* info)
{
	FluidSystem* fs = scratch_system_info(N, steps, info);
	fs->fluid_count = nfluids;
	fs->densList.resize(nfluids);
	fs->fprops.resize(nfluids);
	for (int k = 0; k < nfluids; k++) {
		fs->densList[k].Resize((N + 2) * (N + 2));
		fs->densList[k].Curr() = load(dens[k], N);
		fs->fprops[k].coef[FluidProps::VISC] = coef[k];
		fs->fprops[k].coef[FluidProps::DIFF] = 0;
	}
	vec vx = load(x, N), v0 = load(x0, N);
	fs->LinearSolve1BSO(N, b, vx, v0, aSc, a_fallback, FluidProps::VISC);
	store(x, vx);
	delete fs;
}

void ref_project(int N, float* u, float* v, float* p, float* div, int steps)
{
	FluidSystem* fs = scratch_system(N, steps);
	vec vu = load(u, N), vv = load(v, N), vp = load(p, N), vd = load(div, N);
	fs->Project(N, vu, vv, vp, vd);
	store(u, vu); store(v, vv); store(p, vp); store(div, vd);
	delete fs;
}

} // extern "C"
