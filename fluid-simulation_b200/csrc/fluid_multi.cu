// ---------------------------------------------------------------------------
// fluid_multi: several GPUs from ONE plain-C host process.
//
// The same row-slab contexts as the one-process-per-GPU (NCCL) arrangement, but
// the halo transport is native: peer copies (cudaMemcpyPeerAsync over NVLink)
// ordered purely by CUDA events.  Each slab is stepped by its own host thread
// (the step functions are blocking enqueue loops); at an exchange point a thread
//   1. records "my owned rows are final" on its stream,
//   2. meets the others at a host barrier (so that every such event exists),
//   3. makes its stream wait for its neighbours' events and pulls their boundary
//      rows into its halo rows,
//   4. records "I have read my neighbours' rows", meets the barrier again and
//      makes its stream wait for the neighbours' versions of that event before
//      its next kernel may overwrite the rows they were reading.
// No stream is ever drained and the host never waits for the device: the step
// stays one asynchronous stream of work per GPU.  No counterpart in the
// reference (single process, single thread, no GPU).
// ---------------------------------------------------------------------------
#include "ctx.h"

#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>

namespace {
class HostBarrier {
public:
	explicit HostBarrier(int n) : n_(n) {}
	// returns false when the barrier has been broken by a failing participant
	bool wait()
	{
		std::unique_lock<std::mutex> lk(mu_);
		if (broken_) return false;
		const unsigned gen = gen_;
		if (++count_ == n_) {
			count_ = 0;
			gen_++;
			cv_.notify_all();
			return true;
		}
		cv_.wait(lk, [&] { return gen_ != gen || broken_; });
		return !broken_;
	}
	void abort()
	{
		std::lock_guard<std::mutex> lk(mu_);
		broken_ = true;
		cv_.notify_all();
	}
	void reset()
	{
		std::lock_guard<std::mutex> lk(mu_);
		broken_ = false;
		count_ = 0;
	}

private:
	std::mutex mu_;
	std::condition_variable cv_;
	int n_, count_ = 0;
	unsigned gen_ = 0;
	bool broken_ = false;
};
} // namespace

struct fluid_multi {
	std::vector<fluid_ctx*> slab;
	std::vector<cudaEvent_t> ready, done;
	HostBarrier* barrier = nullptr;
	struct Link { fluid_multi* m; int k; };
	std::vector<Link> links;
	int N = 0, rows = 0;
};

// the exchange callback of slab k (runs on slab k's host thread, inside its step call)
static int multi_exchange(void* user, int nfields, const int* fields)
{
	fluid_multi::Link* ln = (fluid_multi::Link*) user;
	fluid_multi* m = ln->m;
	const int k = ln->k, n = (int) m->slab.size();
	fluid_ctx* c = m->slab[k];
	if (cudaEventRecord(m->ready[k], c->stream) != cudaSuccess) return 10;
	if (!m->barrier->wait()) return 11;
	for (int nb : { k - 1, k + 1 })
		if (nb >= 0 && nb < n && cudaStreamWaitEvent(c->stream, m->ready[nb], 0) != cudaSuccess) return 12;
	for (int q = 0; q < nfields; q++) {
		float *sd, *rd, *su, *ru;
		size_t cnt;
		if (fluid_ctx_halo_regions(c, fields[q], &sd, &rd, &su, &ru, &cnt) != FLUID_OK) return 13;
		if (k > 0) { // my lower halo <- lower neighbour's top owned rows
			float *nsd, *nrd, *nsu, *nru;
			size_t ncnt;
			if (fluid_ctx_halo_regions(m->slab[k - 1], fields[q], &nsd, &nrd, &nsu, &nru, &ncnt) != FLUID_OK) return 13;
			if (cudaSetDevice(c->device) != cudaSuccess) return 14; // halo_regions bound the neighbour's device
			if (cudaMemcpyPeerAsync(rd, c->device, nsu, m->slab[k - 1]->device, cnt * sizeof(float), c->stream) != cudaSuccess) return 15;
		}
		if (k + 1 < n) { // my upper halo <- upper neighbour's bottom owned rows
			float *nsd, *nrd, *nsu, *nru;
			size_t ncnt;
			if (fluid_ctx_halo_regions(m->slab[k + 1], fields[q], &nsd, &nrd, &nsu, &nru, &ncnt) != FLUID_OK) return 13;
			if (cudaSetDevice(c->device) != cudaSuccess) return 14;
			if (cudaMemcpyPeerAsync(ru, c->device, nsd, m->slab[k + 1]->device, cnt * sizeof(float), c->stream) != cudaSuccess) return 15;
		}
	}
	if (cudaEventRecord(m->done[k], c->stream) != cudaSuccess) return 16;
	if (!m->barrier->wait()) return 17;
	for (int nb : { k - 1, k + 1 })
		if (nb >= 0 && nb < n && cudaStreamWaitEvent(c->stream, m->done[nb], 0) != cudaSuccess) return 18;
	return 0;
}

extern "C" int fluid_multi_destroy(fluid_multi* m)
{
	if (!m) return FLUID_OK;
	for (size_t k = 0; k < m->slab.size(); k++) {
		if (m->slab[k]) {
			cudaSetDevice(m->slab[k]->device);
			cudaStreamSynchronize(m->slab[k]->stream);
		}
	}
	for (size_t k = 0; k < m->ready.size(); k++) {
		if (k < m->slab.size() && m->slab[k]) cudaSetDevice(m->slab[k]->device);
		if (m->ready[k]) cudaEventDestroy(m->ready[k]);
		if (m->done[k]) cudaEventDestroy(m->done[k]);
	}
	for (auto* c : m->slab) fluid_ctx_destroy(c);
	delete m->barrier;
	delete m;
	return FLUID_OK;
}

extern "C" int fluid_multi_create(const fluid_params* p, int ndev, const int* devices, fluid_multi** out)
{
	if (!p || !out || !devices || ndev < 1) return fail(FLUID_ERR_INVALID, "bad arguments");
	*out = nullptr;
	if (p->slab_rows < 1) return fail(FLUID_ERR_INVALID, "fluid_multi_create: set slab_rows (rows per device); the global grid is N x slab_rows*ndev");
	fluid_multi* m = new (std::nothrow) fluid_multi();
	if (!m) return fail(FLUID_ERR_NOMEM, "host allocation failed");
	m->N = p->N;
	m->rows = p->slab_rows;
	m->barrier = new HostBarrier(ndev);
	m->slab.assign(ndev, nullptr);
	m->ready.assign(ndev, nullptr);
	m->done.assign(ndev, nullptr);
	m->links.resize(ndev);
	int rc = FLUID_OK;
	for (int k = 0; k < ndev && rc == FLUID_OK; k++) {
		fluid_params q = *p;
		q.device = devices[k];
		q.slab_index = k;
		q.slab_count = ndev;
		q.gs_order = FLUID_GS_REDBLACK;
		if (ndev == 1) { // a single device: one whole grid of N x slab_rows
			q.slab_count = 0;
			q.slab_index = q.slab_halo = 0;
		}
		rc = fluid_ctx_create(&q, &m->slab[k]);
		if (rc != FLUID_OK) break;
		if (cudaEventCreateWithFlags(&m->ready[k], cudaEventDisableTiming) != cudaSuccess ||
		    cudaEventCreateWithFlags(&m->done[k], cudaEventDisableTiming) != cudaSuccess) {
			rc = fail(FLUID_ERR_CUDA, "cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
			break;
		}
		m->links[k] = { m, k };
		m->slab[k]->exchange = multi_exchange;
		m->slab[k]->exchange_user = &m->links[k];
		m->slab[k]->exchange_stream_ordered = true;
	}
	// direct peer access between neighbours (NVLink); without it the copies are staged by the driver
	for (int k = 0; k + 1 < ndev && rc == FLUID_OK; k++) {
		const int a = devices[k], b = devices[k + 1];
		if (a == b) continue;
		int can = 0;
		if (cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) {
			cudaSetDevice(a);
			cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
			if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = fail(FLUID_ERR_CUDA, "enabling peer access %d->%d: %s", a, b, cudaGetErrorString(e));
			cudaSetDevice(b);
			e = cudaDeviceEnablePeerAccess(a, 0);
			if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = fail(FLUID_ERR_CUDA, "enabling peer access %d->%d: %s", b, a, cudaGetErrorString(e));
			cudaGetLastError();
		}
	}
	if (rc != FLUID_OK) {
		char keep[sizeof t_err];
		memcpy(keep, t_err, sizeof keep);
		fluid_multi_destroy(m);
		memcpy(t_err, keep, sizeof keep);
		return rc;
	}
	*out = m;
	return FLUID_OK;
}

#define CHECK_MULTI(m)                                                                                            \
	do {                                                                                                      \
		if (!(m)) return fail(FLUID_ERR_INVALID, "null multi-GPU handle");                                \
	} while (0)

extern "C" int fluid_multi_count(fluid_multi* m) { return m ? (int) m->slab.size() : 0; }

extern "C" int fluid_multi_set_sources(fluid_multi* m, int n, const fluid_source* src)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_set_sources(c, -1, n, src)); // global rows; each slab applies what it stores
	return FLUID_OK;
}

extern "C" int fluid_multi_upload(fluid_multi* m, int field, const float* host_global)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_upload(c, field, 0, host_global));
	return FLUID_OK;
}

extern "C" int fluid_multi_download(fluid_multi* m, int field, float* host_global)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_download(c, field, 0, host_global));
	return FLUID_OK;
}

// fluid_ctx_step on every slab, one host thread per slab; returns when all work is ENQUEUED.
extern "C" int fluid_multi_step(fluid_multi* m, float dt, int nsteps)
{
	CHECK_MULTI(m);
	const int n = (int) m->slab.size();
	std::vector<int> rcs(n, FLUID_OK);
	std::vector<std::string> errs(n);
	m->barrier->reset();
	auto work = [&](int k) {
		rcs[k] = fluid_ctx_step(m->slab[k], dt, nsteps);
		if (rcs[k] != FLUID_OK) {
			errs[k] = t_err; // thread-local message of the worker
			m->barrier->abort();
		}
	};
	std::vector<std::thread> th;
	for (int k = 1; k < n; k++) th.emplace_back(work, k);
	work(0);
	for (auto& t : th) t.join();
	for (int k = 0; k < n; k++)
		if (rcs[k] != FLUID_OK && errs[k].find("callback failed with code 11") == std::string::npos &&
		    errs[k].find("callback failed with code 17") == std::string::npos)
			return fail(rcs[k], "slab %d: %s", k, errs[k].c_str());
	for (int k = 0; k < n; k++)
		if (rcs[k] != FLUID_OK) return fail(rcs[k], "slab %d: %s", k, errs[k].c_str());
	return FLUID_OK;
}

extern "C" int fluid_multi_sync(fluid_multi* m)
{
	CHECK_MULTI(m);
	for (auto* c : m->slab) FL_TRY(fluid_ctx_sync(c));
	return FLUID_OK;
}

// Device time of the last fluid_multi_step: the slowest slab (max over devices), in ms.
extern "C" int fluid_multi_last_step_ms(fluid_multi* m, float* ms)
{
	CHECK_MULTI(m);
	if (!ms) return fail(FLUID_ERR_INVALID, "null pointer");
	float worst = 0;
	for (auto* c : m->slab) {
		float t = 0;
		FL_TRY(fluid_ctx_last_step_ms(c, &t));
		worst = std::max(worst, t);
	}
	*ms = worst;
	return FLUID_OK;
}
