// Gradient exchange fused into the backward (no reference counterpart: MOSS trains on one GPU).
//
// Data-parallel ranks render different views of the same Gaussians and their parameter gradients add.  Instead of
// "all ranks wait, then PULL each other's gradients" (peer_reduce.cuh: barrier, remote loads -- NVLink round trips --,
// remote stores, barrier) every rank PUSHES: each element of its packed gradient is stored into the staging area of the
// rank that owns it, a reduce-scatter made of posted NVLink stores that needs no barrier in front of it.
//   * a batch of views: the view-sum kernel (sum_views_kernel) is the producer and its coalesced 16-byte stores go
//     straight to the owners while it is still summing (route());
//   * one view: the backward preprocess writes 4..16-byte pieces per thread at 12..192-byte strides, which NVLink
//     turns into one small packet each (measured: K7 13 -> 31 us with routed stores at 2 ranks), so it keeps writing
//     locally and the gather kernel below opens with a coalesced push of the L2-resident rows instead.
// What is left behind the producer is ONE short kernel:
//   [push my rows]  signal "my rows are out" to every owner -> wait until all ranks' rows have arrived in my staging area
//   sum the `world` staged copies of my slice in rank order (local reads; bit-reproducible)
//   store the result into slice `rank` of EVERY rank's gradient buffer (one multimem.st through the switch, or
//   `world` peer stores), signal "slice delivered" -> wait for the other owners' slices.
// Ownership is by position in the packed buffer: element j of the summed prefix belongs to rank j / S
// (S = slice length, a multiple of 4 floats so a 16-byte store never straddles two owners).
// Flags live in the peers' signal pads exactly like peer_reduce.cuh (slot ((phase * blocks + block) * world + writer),
// value = a monotonically increasing epoch kept on the device, so the call is CUDA-graph capturable).
#pragma once
#include "common.cuh"
#include "peer_reduce.cuh"

namespace moss {

// Where a producer stores element `off` of the packed gradient.  numel == 0: push disabled (everything local).
struct PushCfg {
	const float* base;               // this rank's packed gradient buffer
	uint32_t numel;                  // floats of the summed prefix (multiple of 4)
	uint32_t S;                      // floats per owner slice (multiple of 4)
	float* stage_mine[PEER_MAX_WORLD];   // stage_mine[o] = rank o's staging area + my_rank * S, as mapped in this process
};
__device__ __forceinline__ float* route(const PushCfg& pc, float* p)
{
	if (pc.numel == 0u) return p;
	const size_t off = (size_t)(p - pc.base);          // wraps to a huge value for pointers in front of the buffer
	if (off >= (size_t)pc.numel) return p;             // blend-only outputs (colours, conics, per-view rows) stay local
	const uint32_t o32 = (uint32_t)off;
	const uint32_t owner = o32 / pc.S;
	return pc.stage_mine[owner] + (o32 - owner * pc.S);
}
inline PushCfg no_push()
{
	PushCfg pc;
	pc.base = nullptr; pc.numel = 0; pc.S = 4;
	for (int i = 0; i < PEER_MAX_WORLD; i++) pc.stage_mine[i] = nullptr;
	return pc;
}

struct GatherArgs {
	const float* push_src;           // my packed gradient if this kernel does the push itself, null if the producer did
	float* stage_mine[PEER_MAX_WORLD];   // stage_mine[o] = rank o's staging area + my_rank * S
	float* grad[PEER_MAX_WORLD];     // every rank's packed gradient buffer, as mapped here
	uint32_t* pads[PEER_MAX_WORLD];  // signal pads
	const float* stage;              // MY staging area: [world][S]
	float* mc_grad;                  // multicast mapping of the gradient buffer, or null
	uint32_t* err;                   // [0] timeout, [1] epoch, [2] block counter
	uint32_t numel, S;
	int rank, world;
};
constexpr int GATHER_THREADS = 256;
constexpr int GATHER_ILP = 8;          // independent 16-byte loads in flight per thread (push and sum)

__device__ __forceinline__ void flag_exchange(uint32_t* const* pads, int rank, int world, int phase, uint32_t epoch, uint32_t* err)
{
	// every thread's stores of this block must be visible system-wide before the flag: the block syncs, then the
	// signalling threads fence at system scope (cumulative over what the barrier made visible to them)
	__syncthreads();
	if ((int)threadIdx.x < world) {
		const int p = (int)threadIdx.x;
		const size_t slot = (size_t)(phase * (int)gridDim.x + (int)blockIdx.x) * world;
		__threadfence_system();
		st_relaxed_sys(pads[p] + slot + rank, epoch);
		const uint32_t* mine = pads[rank] + slot + p;
		const long long t0 = clock64();
		while ((int)(ld_relaxed_sys(mine) - epoch) < 0) {
			if (clock64() - t0 > 20000000000ll) { atomicExch(err, 1u); break; }    // ~10 s at 2 GHz: a peer died
		}
		__threadfence_system();
	}
	__syncthreads();
}

__global__ void __launch_bounds__(GATHER_THREADS)
push_gather_kernel(const __grid_constant__ GatherArgs a)
{
	const uint32_t epoch = ld_relaxed_sys(a.err + 1) + 1u;
	__shared__ uint32_t* s_pads[PEER_MAX_WORLD];
	if ((int)threadIdx.x < a.world) s_pads[threadIdx.x] = a.pads[threadIdx.x];
	const uint32_t stride = gridDim.x * GATHER_THREADS;
	const uint32_t S4 = a.S >> 2;
	// The exchange is a chain of latencies (two flag phases, two NVLink hops); what can be trimmed is the time the data
	// passes take, so every thread issues all its loads before its first store and the grid is as wide as the flag
	// slots allow: at ZJU size (4.9 MB of gradient, 8 ranks) a thread moves 16 + 16 float4 in two bursts.
	const uint32_t rounds = (S4 + stride - 1) / stride;                 // float4 per thread per owner slice
	if (a.push_src) {
		// Block b pushes, for every owner o, exactly the elements block b of rank o will sum (same index pattern), so a
		// per-block flag is all the synchronisation the data needs.  Work item w = (round, owner), owner fastest.
		const float4* src4 = reinterpret_cast<const float4*>(a.push_src);
		const uint32_t items = rounds * (uint32_t)a.world;
		for (uint32_t w0 = 0; w0 < items; w0 += GATHER_ILP) {
			float4 v[GATHER_ILP];
			bool ok[GATHER_ILP];
#pragma unroll
			for (int u = 0; u < GATHER_ILP; u++) {
				const uint32_t w = w0 + u, o = w % (uint32_t)a.world, r = w / (uint32_t)a.world;
				const uint32_t i = r * stride + blockIdx.x * GATHER_THREADS + threadIdx.x;
				const uint32_t olo = min(a.numel, a.S * o), ohi = min(a.numel, olo + a.S);
				ok[u] = w < items && i < ((ohi - olo) >> 2);
				if (ok[u]) v[u] = __ldcg(src4 + (olo >> 2) + i);
			}
#pragma unroll
			for (int u = 0; u < GATHER_ILP; u++) {
				const uint32_t w = w0 + u, o = w % (uint32_t)a.world, r = w / (uint32_t)a.world;
				const uint32_t i = r * stride + blockIdx.x * GATHER_THREADS + threadIdx.x;
				if (ok[u]) st_peer_f4(reinterpret_cast<float4*>(a.stage_mine[o]) + i, v[u]);
			}
		}
	}
	// phase 0: the producers of all ranks (earlier kernels in their streams) have pushed their rows into my staging area
	flag_exchange(s_pads, a.rank, a.world, 0, epoch, a.err);

	const uint32_t lo = min(a.numel, a.S * (uint32_t)a.rank), hi = min(a.numel, lo + a.S);   // my slice of the prefix
	const uint32_t n4 = (hi - lo) >> 2;
	const float4* st4 = reinterpret_cast<const float4*>(a.stage);
	for (uint32_t r = 0; r < rounds; r++) {
		const uint32_t i = r * stride + blockIdx.x * GATHER_THREADS + threadIdx.x;
		if (i >= n4) break;
		float4 x[PEER_MAX_WORLD];
#pragma unroll
		for (int p = 0; p < PEER_MAX_WORLD; p++)
			if (p < a.world) x[p] = __ldcg(st4 + (size_t)p * S4 + i);          // the `world` staged copies: independent loads
		float4 s = x[0];
#pragma unroll
		for (int p = 1; p < PEER_MAX_WORLD; p++)
			if (p < a.world) { s.x += x[p].x; s.y += x[p].y; s.z += x[p].z; s.w += x[p].w; }      // rank order: reproducible
		const size_t e4 = ((size_t)lo >> 2) + i;
		if (a.mc_grad) multimem_st_f4(reinterpret_cast<float4*>(a.mc_grad) + e4, s);
		else {
#pragma unroll
			for (int p = 0; p < PEER_MAX_WORLD; p++)
				if (p < a.world) st_peer_f4(reinterpret_cast<float4*>(a.grad[p]) + e4, s);
		}
	}
	// phase 1: my slice is in everybody's buffer; wait for the other owners' slices
	flag_exchange(s_pads, a.rank, a.world, 1, epoch, a.err);
	if (threadIdx.x == 0) {
		if (atomicAdd(a.err + 2, 1u) == gridDim.x - 1u) { a.err[2] = 0u; __threadfence(); a.err[1] = epoch; }
	}
}

} // namespace moss
