// Native halo transport between row slabs: pushes over NVLink peer memory, ordered by flags.
//
// Measured in round 1 (profiles/r01_bench_8gpu.json): the six halo refreshes of a frame cost 0.74 ms of a
// 6.8 ms frame on 8 GPUs although they move only 16 MB per rank (21 us at NVLink speed) -- each refresh is an
// NCCL send/recv group behind two stream hops, ~120 us of latency during which the SMs idle.  Here a refresh
// is two small kernels on the context's own stream:
//   k_halo_push   copies this slab's boundary rows straight into the NEIGHBOUR's inbox (plain st.global through a
//                 peer pointer: the neighbour's memory mapped by CUDA IPC, or directly addressable when both slabs
//                 live in one process), then -- last block, after a system-wide fence -- stores the refresh's
//                 sequence number into the neighbour's flag word;
//   k_halo_pull   waits until this slab's own flag words show that sequence number (its neighbours' pushes have
//                 landed), then copies its inboxes into its halo rows.
// No host involvement, no collective, no copy engine: ~25 us per refresh, and the two kernels queue behind the
// compute kernels like any others.  Inboxes are double-buffered by the parity of the sequence number: a
// neighbour can push refresh s+2 (same buffer as s) only after its own pull of s+1, which needs this slab's push of
// s+1, which this slab's stream runs after its pull of s.  Waiting kernels of different slabs run on different
// GPUs, so nobody waits for a kernel that shares its device.  No counterpart in the reference.
#include "ctx.h"

#include <stdlib.h>

namespace {

struct PushArgs {
	const float4* src_up[2];   // my last halo+1 owned rows, per field
	const float4* src_down[2]; // my first halo+1 owned rows
	float4* dst_up[2];         // upper neighbour's inbox "from below"
	float4* dst_down[2];       // lower neighbour's inbox "from above"
	unsigned int* flag_up;     // upper neighbour's flag "from below" (nullptr: no such neighbour)
	unsigned int* flag_down;   // lower neighbour's flag "from above"
	int nfields;
	long long n4;              // float4 per field block
	unsigned int seq;
};

struct PullArgs {
	const float4* in_below[2]; // my inbox "from below" -> my lower halo rows
	const float4* in_above[2];
	float4* dst_low[2];
	float4* dst_high[2];
	const unsigned int* flag_below; // nullptr: no lower neighbour
	const unsigned int* flag_above;
	int nfields;
	long long n4;
	unsigned int seq;
};

__global__ void __launch_bounds__(256) k_halo_push(PushArgs a, unsigned int* __restrict__ done_blocks)
{
	const long long stride = (long long) gridDim.x * blockDim.x;
	for (int f = 0; f < a.nfields; f++) {
		for (long long k = (long long) blockIdx.x * blockDim.x + threadIdx.x; k < a.n4; k += stride) {
			if (a.flag_up) a.dst_up[f][k] = a.src_up[f][k];
			if (a.flag_down) a.dst_down[f][k] = a.src_down[f][k];
		}
	}
	// every block's stores must be visible on the neighbour before the flag is: fence, count, last block publishes
	__threadfence_system();
	__syncthreads();
	if (threadIdx.x == 0) {
		const unsigned int n = atomicAdd(done_blocks, 1u) + 1u;
		if (n == gridDim.x) {
			*done_blocks = 0;
			__threadfence_system();
			if (a.flag_up) *(volatile unsigned int*) a.flag_up = a.seq;
			if (a.flag_down) *(volatile unsigned int*) a.flag_down = a.seq;
			__threadfence_system();
		}
	}
}

// A neighbour that never pushes (its process died, the slabs made different sequences of calls) must not hang this
// GPU: after ~4e9 clock cycles (about two seconds) of waiting the kernel gives up, raises bit 1 of *fault and lets the
// step run on with stale halo rows; fluid_ctx_sync reports it.
__device__ __forceinline__ void wait_for_flag(const unsigned int* flag, unsigned int seq, int* fault)
{
	const long long t0 = clock64();
	while ((int) (*(volatile const unsigned int*) flag - seq) < 0) { // sequence numbers only grow; signed difference survives the wrap
		__nanosleep(200);
		if (clock64() - t0 > 4000000000LL) {
			atomicOr(fault, 2);
			break;
		}
	}
}

__global__ void __launch_bounds__(256) k_halo_pull(PullArgs a, int* __restrict__ fault)
{
	if (threadIdx.x == 0) {
		if (a.flag_below) wait_for_flag(a.flag_below, a.seq, fault);
		if (a.flag_above) wait_for_flag(a.flag_above, a.seq, fault);
		__threadfence_system();
	}
	__syncthreads();
	const long long stride = (long long) gridDim.x * blockDim.x;
	for (int f = 0; f < a.nfields; f++) {
		for (long long k = (long long) blockIdx.x * blockDim.x + threadIdx.x; k < a.n4; k += stride) {
			// (the inbox was written by another GPU: read it past L1, which may hold the lines of two refreshes ago)
			if (a.flag_below) a.dst_low[f][k] = __ldcg(a.in_below[f] + k);
			if (a.flag_above) a.dst_high[f][k] = __ldcg(a.in_above[f] + k);
		}
	}
}

constexpr int kPeerFields = 2;  // fields per refresh at most ({u, v})
constexpr int kPeerSlots = 2;   // double buffering
constexpr size_t kFlagBytes = 256;

inline size_t inbox_floats(const fluid_ctx* c) { return (size_t) (c->slab_halo + 1) * c->g.pitch; }
inline size_t block_bytes(const fluid_ctx* c) { return kFlagBytes + 2 * kPeerSlots * kPeerFields * inbox_floats(c) * sizeof(float); }
// layout of a block: [flags: from_below, from_above | pad to 256 B][from_below: slot 0 field 0, field 1, slot 1 ...][from_above: ...]
inline float* inbox(void* block, const fluid_ctx* c, int from_above, int slot, int field)
{
	return (float*) ((char*) block + kFlagBytes) + ((size_t) (from_above * kPeerSlots + slot) * kPeerFields + field) * inbox_floats(c);
}
inline unsigned int* flag(void* block, int from_above) { return (unsigned int*) block + from_above; }

} // namespace

extern "C" int fluid_ctx_peer_halo_local(fluid_ctx* c, void** block, size_t* bytes)
{
	CHECK_CTX(c);
	if (!block || !bytes) return fail(FLUID_ERR_INVALID, "null pointer");
	if (c->slab_count <= 1) return fail(FLUID_ERR_INVALID, "not a row slab");
	if (!c->peer_block) {
		CU_TRY(cudaMalloc(&c->peer_block, block_bytes(c)));
		CU_TRY(cudaMemset(c->peer_block, 0, block_bytes(c)));
		CU_TRY(cudaMalloc(&c->peer_counter, sizeof(unsigned int)));
		CU_TRY(cudaMemset(c->peer_counter, 0, sizeof(unsigned int)));
	}
	*block = c->peer_block;
	*bytes = block_bytes(c);
	return FLUID_OK;
}

extern "C" int fluid_ctx_peer_halo_connect(fluid_ctx* c, void* lower_block, void* upper_block)
{
	CHECK_CTX(c);
	if (!c->peer_block) return fail(FLUID_ERR_INVALID, "call fluid_ctx_peer_halo_local first");
	const bool lo_nb = c->slab_index > 0, hi_nb = c->slab_index < c->slab_count - 1;
	if ((lo_nb && !lower_block) || (hi_nb && !upper_block)) return fail(FLUID_ERR_INVALID, "a neighbour's block is missing");
	c->peer_lower = lo_nb ? lower_block : nullptr;
	c->peer_upper = hi_nb ? upper_block : nullptr;
	c->peer_connected = true;
	c->peer_seq = 0;
	return FLUID_OK;
}

// One refresh of the listed fields (context field ids), enqueued on c->stream.
int peer_halo_exchange(fluid_ctx* c, int n, const int* ids)
{
	if (n > kPeerFields) return fail(FLUID_ERR_CUDA, "internal: %d fields in one halo refresh", n);
	const unsigned int seq = ++c->peer_seq;
	const int slot = (int) (seq & 1u);
	PushArgs pa = {};
	PullArgs pl = {};
	pa.nfields = pl.nfields = n;
	pa.n4 = pl.n4 = (long long) (inbox_floats(c) / 4);
	pa.seq = pl.seq = seq;
	pa.flag_up = c->peer_upper ? flag(c->peer_upper, 0) : nullptr;   // I am below my upper neighbour
	pa.flag_down = c->peer_lower ? flag(c->peer_lower, 1) : nullptr; // ... and above my lower one
	pl.flag_below = c->peer_lower ? flag(c->peer_block, 0) : nullptr;
	pl.flag_above = c->peer_upper ? flag(c->peer_block, 1) : nullptr;
	for (int q = 0; q < n; q++) {
		float *sd, *rd, *su, *ru;
		size_t cnt;
		FL_TRY(fluid_ctx_halo_regions(c, ids[q], &sd, &rd, &su, &ru, &cnt));
		pa.src_up[q] = (const float4*) su;
		pa.src_down[q] = (const float4*) sd;
		pa.dst_up[q] = c->peer_upper ? (float4*) inbox(c->peer_upper, c, 0, slot, q) : nullptr;
		pa.dst_down[q] = c->peer_lower ? (float4*) inbox(c->peer_lower, c, 1, slot, q) : nullptr;
		pl.in_below[q] = (const float4*) inbox(c->peer_block, c, 0, slot, q);
		pl.in_above[q] = (const float4*) inbox(c->peer_block, c, 1, slot, q);
		pl.dst_low[q] = (float4*) rd;
		pl.dst_high[q] = (float4*) ru;
	}
	static const int env_blocks = getenv("FLUID_PEER_BLOCKS") ? atoi(getenv("FLUID_PEER_BLOCKS")) : 0;
	const int cap = env_blocks > 0 ? env_blocks : 2 * c->num_sms;
	const int blocks = (int) std::min<long long>((pa.n4 + 255) / 256, (long long) cap);
	LAUNCH(k_halo_push, blocks, 256, 0, c->stream, pa, c->peer_counter);
	LAUNCH(k_halo_pull, blocks, 256, 0, c->stream, pl, c->d_oob);
	CU_TRY(cudaGetLastError());
	return FLUID_OK;
}

// CUDA IPC for hosts that run one process per GPU: 64-byte handles of device allocations.
extern "C" int fluid_ipc_export(void* device_ptr, unsigned char handle[64])
{
	if (!device_ptr || !handle) return fail(FLUID_ERR_INVALID, "null pointer");
	static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
	cudaIpcMemHandle_t h;
	CU_TRY(cudaIpcGetMemHandle(&h, device_ptr));
	memcpy(handle, &h, 64);
	return FLUID_OK;
}

extern "C" int fluid_ipc_import(const unsigned char handle[64], void** device_ptr)
{
	if (!device_ptr || !handle) return fail(FLUID_ERR_INVALID, "null pointer");
	cudaIpcMemHandle_t h;
	memcpy(&h, handle, 64);
	CU_TRY(cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
	return FLUID_OK;
}

extern "C" int fluid_ipc_close(void* device_ptr)
{
	if (device_ptr) CU_TRY(cudaIpcCloseMemHandle(device_ptr));
	return FLUID_OK;
}
