// Nearest SMPL vertex of every Gaussian.
//
// MOSS looks every Gaussian up against the big-pose SMPL vertices on EVERY render, through the third-party
// knn_cuda wheel: `_, vert_ids = self.knn(smpl_pts.float(), query_pts.float())` with KNN(k=1, transpose_mode=True)
// (scene/gaussian_model.py:85, :827; the distance output is what :657 reads).  That wheel is not part of the
// reference tree; its published algorithm for k=1 is the exhaustive search restated here: the vertex with the smallest
// squared Euclidean distance, the lowest index winning ties.
//
// A block owns 32 Gaussians (one per lane) and stages the vertices through shared memory in tiles; its 8 warps each
// scan a different eighth of every tile for the same 32 Gaussians (every lane of a warp reads the same vertex: a
// broadcast, no bank conflicts), so 20 k Gaussians put ~34 warps on every SM instead of 4 (one thread per Gaussian
// scanning all 6890 vertices left the schedulers idle: 124 us, 27 % of the issue rate).  Four independent
// (distance, index) minima per thread keep the compare/select chain off the critical path; all partial minima are
// merged lexicographically -- smaller distance, then smaller index -- so the result is the lowest index among equal
// distances whatever the split.  Distances are formed from the coordinate differences (not from |a|^2+|b|^2-2ab,
// which cancels catastrophically for a point ON the surface) in a fixed, explicitly rounded order.
// 20 k Gaussians x 6890 vertices = 138 M candidate pairs, all from shared memory: instruction-issue bound; HBM sees
// P*12 + V*12 bytes in and P*4 out, the vertex tiles are re-read from L2 once per block.
#pragma once
#include "common.cuh"

namespace moss {

constexpr int NEAR_WARPS = 8;
constexpr int NEAR_THREADS = 32 * NEAR_WARPS;
constexpr int NEAR_TILE = 1024;                       // vertices per shared-memory tile (16 KB as float4)
constexpr int NEAR_SLICE = NEAR_TILE / NEAR_WARPS;    // vertices of a tile one warp scans

__device__ __forceinline__ void nearer(float& bd, int& bi, float d, int i)
{
	const bool lt = (d < bd) || (d == bd && i < bi);   // lexicographic: merges are order-independent
	bd = lt ? d : bd;
	bi = lt ? i : bi;
}

__global__ void __launch_bounds__(NEAR_THREADS)
nearest_vertex_kernel(int P, const float* __restrict__ xyz, int V, const float* __restrict__ verts,
	int* __restrict__ vert_ids, float* __restrict__ dist)
{
	__shared__ float4 tile[NEAR_TILE];
	__shared__ float s_d[NEAR_WARPS][32];
	__shared__ int s_i[NEAR_WARPS][32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int idx = blockIdx.x * 32 + lane;
	float qx = 0.f, qy = 0.f, qz = 0.f;
	if (idx < P) { qx = xyz[3 * idx]; qy = xyz[3 * idx + 1]; qz = xyz[3 * idx + 2]; }
	const float INF = __int_as_float(0x7f800000);
	float bd[4] = {INF, INF, INF, INF};
	int bi[4] = {0x7fffffff, 0x7fffffff, 0x7fffffff, 0x7fffffff};
	for (int base = 0; base < V; base += NEAR_TILE) {
		const int m = min(NEAR_TILE, V - base);
		__syncthreads();
		for (int j = threadIdx.x; j < NEAR_TILE; j += NEAR_THREADS) {
			float4 v = make_float4(INF, INF, INF, 0.f);      // padding never wins: (q - inf)^2 = inf
			if (j < m) { const float* s = verts + 3 * (size_t)(base + j); v = make_float4(s[0], s[1], s[2], 0.f); }
			tile[j] = v;
		}
		__syncthreads();
		const int lo = warp * NEAR_SLICE;
		if (lo < m) {
#pragma unroll 4
			for (int j = lo; j < lo + NEAR_SLICE; j += 4) {
#pragma unroll
				for (int u = 0; u < 4; u++) {
					const float4 v = tile[j + u];
					const float dx = __fsub_rn(qx, v.x), dy = __fsub_rn(qy, v.y), dz = __fsub_rn(qz, v.z);
					const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
					// ascending scan inside one accumulator: a strict < keeps the lowest index of equal distances
					const bool lt = d2 < bd[u];
					bd[u] = lt ? d2 : bd[u];
					bi[u] = lt ? base + j + u : bi[u];
				}
			}
		}
	}
#pragma unroll
	for (int u = 1; u < 4; u++) nearer(bd[0], bi[0], bd[u], bi[u]);
	s_d[warp][lane] = bd[0];
	s_i[warp][lane] = bi[0];
	__syncthreads();
	if (warp == 0 && idx < P) {
		float d = s_d[0][lane];
		int i = s_i[0][lane];
#pragma unroll
		for (int w = 1; w < NEAR_WARPS; w++) nearer(d, i, s_d[w][lane], s_i[w][lane]);
		vert_ids[idx] = i < V ? i : 0;                   // a NaN query compares false against everything
		if (dist) dist[idx] = __fsqrt_rn(d);
	}
}

} // namespace moss
