// Packed-gradient all-reduce over NVLink peer memory: the one exchange step of the
// data-parallel path.  Every rank's backward writes its packed parameter gradient into a SYMMETRIC buffer (same
// allocation on every GPU, all of them mapped into every process); this kernel then does a two-shot all-reduce with
// plain loads/stores on the mapped peer pointers:
//   barrier  -> every peer's gradient is complete (its backward precedes this kernel in stream order)
//   reduce   -> rank r sums slice r of all `world` buffers in rank order (deterministic) and stores the result into
//               slice r of EVERY buffer; slice r is read by rank r only, so no buffer is written while it is read
//   barrier  -> every slice of the local buffer has arrived
// 4.9 MB at ZJU shape: (world-1)/world of it crosses NVLink once in each direction, versus NCCL's ~50-70 us launch +
// protocol latency at this size, which is a third of the 0.2 ms step.
// Flags live in the peers' signal pads: slot ((phase * gridDim.x + block) * world + writer), value = the call's
// epoch (monotonic, never reset).  Spins give up after ~10 s and raise the error word instead of hanging the GPU.
#pragma once
#include "common.cuh"

namespace moss {

constexpr int PEER_MAX_BLOCKS = 128;   // upper bound; the launch uses what the signal pads can hold
constexpr int PEER_ILP = 4;            // float4 per thread in flight: an NVLink round trip is ~2 us, so the loop must not serialise them
constexpr int PEER_MAX_WORLD = 8;
constexpr int PEER_THREADS = 256;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v)
{
	asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p)
{
	uint32_t v;
	asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p)
{
	float4 v;
	asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_peer_f4(float4* p, float4 v)
{
	asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// NVLS: one instruction loads the SUM over all ranks of a multicast address (reduced inside the NVSwitch) / stores to
// all ranks at once.  Used when the symmetric buffer has a multicast mapping.
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float4* mc)
{
	float4 v;
	asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
		: "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
	return v;
}
__device__ __forceinline__ void multimem_st_f4(float4* mc, float4 v)
{
	asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v)
{
	asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p)
{
	uint32_t v;
	asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}

// One flag exchange between block b of this rank and block b of every peer.  Relaxed flag accesses bracketed by
// system-scope fences (fence -> store ... load -> fence): polling with acquire loads and signalling with release
// stores made every poll / signal a full ordering point and cost several microseconds per barrier.
// `publish`: the block wrote data that the peers will read after the barrier.
__device__ __forceinline__ void peer_barrier(uint32_t* const* pads, int rank, int world, int phase, uint32_t epoch, uint32_t* err,
	bool publish)
{
	__syncthreads();
	if ((int)threadIdx.x < world) {
		const int p = (int)threadIdx.x;
		const size_t slot = (size_t)(phase * (int)gridDim.x + (int)blockIdx.x) * world;
		if (publish) __threadfence_system();
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

__global__ void __launch_bounds__(PEER_THREADS)
peer_allreduce_kernel(float* const* bufs, uint32_t* const* pads, int rank, int world, size_t offset, size_t numel,
	uint32_t epoch, uint32_t* err, float* mc)
{
	// device-side epoch (graph-capturable): err[1] holds the number of completed exchanges; the LAST block to finish
	// bumps it, and by then every block of this launch has read it
	const bool dev_epoch = epoch == 0;
	if (dev_epoch) epoch = ld_relaxed_sys(err + 1) + 1u;
	__shared__ float* s_buf[PEER_MAX_WORLD];
	if ((int)threadIdx.x < world) s_buf[threadIdx.x] = bufs[threadIdx.x] + offset;
	peer_barrier(pads, rank, world, 0, epoch, err, false);     // the peers' gradients were written by earlier kernels

	const size_t n4 = numel / 4;
	const size_t per = (n4 + world - 1) / world;                 // float4s per rank slice
	const size_t lo = min(n4, per * rank), hi = min(n4, lo + per);
	const size_t stride = (size_t)gridDim.x * PEER_THREADS;
	if (mc) {
		// in-switch reduction: (1/world) of the buffer in, the same out, per GPU, whatever the world size
		float4* m4 = reinterpret_cast<float4*>(mc + offset);
		for (size_t i0 = lo + (size_t)blockIdx.x * PEER_THREADS + threadIdx.x; i0 < hi; i0 += stride * PEER_ILP) {
			float4 v[PEER_ILP];
#pragma unroll
			for (int u = 0; u < PEER_ILP; u++)
				if (i0 + u * stride < hi) v[u] = multimem_ld_reduce_f4(m4 + i0 + u * stride);
#pragma unroll
			for (int u = 0; u < PEER_ILP; u++)
				if (i0 + u * stride < hi) multimem_st_f4(m4 + i0 + u * stride, v[u]);
		}
	} else
	for (size_t i0 = lo + (size_t)blockIdx.x * PEER_THREADS + threadIdx.x; i0 < hi; i0 += stride * PEER_ILP) {
		float4 v[PEER_ILP][PEER_MAX_WORLD];
#pragma unroll
		for (int u = 0; u < PEER_ILP; u++) {
			const size_t i = i0 + u * stride;
#pragma unroll
			for (int p = 0; p < PEER_MAX_WORLD; p++)
				if (p < world && i < hi) v[u][p] = ld_peer_f4(reinterpret_cast<const float4*>(s_buf[p]) + i);
		}
#pragma unroll
		for (int u = 0; u < PEER_ILP; u++) {
			const size_t i = i0 + u * stride;
			if (i < hi) {
				float4 s = v[u][0];
#pragma unroll
				for (int p = 1; p < PEER_MAX_WORLD; p++)
					if (p < world) { s.x += v[u][p].x; s.y += v[u][p].y; s.z += v[u][p].z; s.w += v[u][p].w; }
#pragma unroll
				for (int p = 0; p < PEER_MAX_WORLD; p++)
					if (p < world) st_peer_f4(reinterpret_cast<float4*>(s_buf[p]) + i, s);
			}
		}
	}
	// scalar tail (numel % 4) by the last rank
	if (rank == world - 1 && blockIdx.x == 0 && threadIdx.x < (numel & 3)) {
		const size_t i = n4 * 4 + threadIdx.x;
		float s = 0.f;
		for (int p = 0; p < world; p++) s += *((volatile float*)(s_buf[p] + i));
		for (int p = 0; p < world; p++) *((volatile float*)(s_buf[p] + i)) = s;
	}
	peer_barrier(pads, rank, world, 1, epoch, err, true);
	if (dev_epoch && threadIdx.x == 0) {
		if (atomicAdd(err + 2, 1u) == gridDim.x - 1u) { err[2] = 0u; __threadfence(); err[1] = epoch; }
	}
}

} // namespace moss
