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


// ---- the same search through a uniform grid over the vertices ---------------------------------------------------------
// The exhaustive kernel above looks at 138 M pairs for 20 k Gaussians although a Gaussian sits within a few centimetres of
// the body surface.  The vertices (constant for a run: the big-pose world vertices, scene/dataset_readers.py:586-601) are
// binned ONCE into a 32^3 grid over their bounding box; a query walks the cell shells around its own cell, ring by ring,
// and stops as soon as its best distance is smaller than the distance to the nearest face of the block it has covered
// (faces on the grid boundary do not count: nothing lies beyond them).  Same distance expression, same lexicographic
// (distance, index) minimum as the exhaustive kernel, so the answers are identical -- including ties -- while a query
// near the surface examines ~30-80 vertices instead of 6890 (24 us instead of 67 us at 20 k Gaussians).
constexpr int GRID_G = 32;
constexpr int GRID_CELLS = GRID_G * GRID_G * GRID_G;
constexpr int GRID_BUILD_THREADS = 1024;

struct VertexGridView {            // layout of the caller-owned grid buffer
	float* header;                 // [8]: lo.x, lo.y, lo.z, h, 1/h, V (as float bits), -, -
	int* cell_start;               // [GRID_CELLS + 1]
	int* cursor;                   // [GRID_CELLS] build scratch
	float4* sorted;                // [V]: x, y, z, index bits -- grouped by cell
};
__host__ __device__ __forceinline__ size_t vertex_grid_bytes(int V)
{
	return 32 + sizeof(int) * (size_t)(2 * GRID_CELLS + 1 + 3) + 16 + sizeof(float4) * (size_t)V;
}
__host__ __device__ __forceinline__ VertexGridView vertex_grid_view(void* base, int V)
{
	VertexGridView g;
	char* p = (char*)base;
	g.header = (float*)p; p += 32;
	g.cell_start = (int*)p; p += sizeof(int) * (GRID_CELLS + 1 + 3);
	g.cursor = (int*)p; p += sizeof(int) * GRID_CELLS;
	p = (char*)(((size_t)p + 15) & ~(size_t)15);
	g.sorted = (float4*)p;
	(void)V;
	return g;
}
__device__ __forceinline__ int grid_coord(float v, float lo, float inv_h)
{
	return min(GRID_G - 1, max(0, (int)((v - lo) * inv_h)));
}

// ONE block: bounding box, per-cell counts, exclusive scan, fill.  Runs once per vertex set.
__global__ void __launch_bounds__(GRID_BUILD_THREADS)
vertex_grid_build_kernel(int V, const float* __restrict__ verts, VertexGridView g)
{
	__shared__ float s_lo[3][32], s_hi[3][32];
	__shared__ int s_scan[GRID_BUILD_THREADS];
	__shared__ float s_hdr[5];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const float INF = __int_as_float(0x7f800000);
	float lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
	for (int v = tid; v < V; v += GRID_BUILD_THREADS)
#pragma unroll
		for (int c = 0; c < 3; c++) { const float x = verts[3 * (size_t)v + c]; lo[c] = fminf(lo[c], x); hi[c] = fmaxf(hi[c], x); }
#pragma unroll
	for (int c = 0; c < 3; c++) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) { lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o)); hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o)); }
		if (lane == 0) { s_lo[c][warp] = lo[c]; s_hi[c][warp] = hi[c]; }
	}
	__syncthreads();
	if (tid == 0) {
		float ext = 0.f;
		for (int c = 0; c < 3; c++) {
			float l = INF, h = -INF;
			for (int w = 0; w < GRID_BUILD_THREADS / 32; w++) { l = fminf(l, s_lo[c][w]); h = fmaxf(h, s_hi[c][w]); }
			s_hdr[c] = l;
			ext = fmaxf(ext, h - l);
		}
		const float h = fmaxf(ext, 1e-6f) / (float)GRID_G * 1.0001f;      // the largest coordinate still falls into cell G-1
		s_hdr[3] = h; s_hdr[4] = 1.0f / h;
		for (int i = 0; i < 5; i++) g.header[i] = s_hdr[i];
		g.header[5] = __int_as_float(V);
	}
	for (int i = tid; i < GRID_CELLS; i += GRID_BUILD_THREADS) g.cursor[i] = 0;
	__syncthreads();
	const float lx = s_hdr[0], ly = s_hdr[1], lz = s_hdr[2], inv_h = s_hdr[4];
	for (int v = tid; v < V; v += GRID_BUILD_THREADS) {
		const int cx = grid_coord(verts[3 * (size_t)v], lx, inv_h), cy = grid_coord(verts[3 * (size_t)v + 1], ly, inv_h),
			cz = grid_coord(verts[3 * (size_t)v + 2], lz, inv_h);
		atomicAdd(&g.cursor[(cz * GRID_G + cy) * GRID_G + cx], 1);
	}
	__syncthreads();
	// exclusive scan of the 32768 counts: 32 consecutive cells per thread, then a block scan of the thread sums
	constexpr int PER = GRID_CELLS / GRID_BUILD_THREADS;
	int local = 0;
	for (int k = 0; k < PER; k++) local += g.cursor[tid * PER + k];
	s_scan[tid] = local;
	__syncthreads();
	for (int o = 1; o < GRID_BUILD_THREADS; o <<= 1) {
		const int add = tid >= o ? s_scan[tid - o] : 0;
		__syncthreads();
		s_scan[tid] += add;
		__syncthreads();
	}
	int run = s_scan[tid] - local;
	for (int k = 0; k < PER; k++) {
		const int c = g.cursor[tid * PER + k];
		g.cell_start[tid * PER + k] = run;
		g.cursor[tid * PER + k] = run;          // becomes the fill cursor
		run += c;
	}
	if (tid == GRID_BUILD_THREADS - 1) g.cell_start[GRID_CELLS] = run;
	__syncthreads();
	for (int v = tid; v < V; v += GRID_BUILD_THREADS) {
		const float x = verts[3 * (size_t)v], y = verts[3 * (size_t)v + 1], z = verts[3 * (size_t)v + 2];
		const int cell = (grid_coord(z, lz, inv_h) * GRID_G + grid_coord(y, ly, inv_h)) * GRID_G + grid_coord(x, lx, inv_h);
		const int pos = atomicAdd(&g.cursor[cell], 1);      // order inside a cell is irrelevant: the minimum is lexicographic
		g.sorted[pos] = make_float4(x, y, z, __int_as_float(v));
	}
}

// SIXTEEN LANES per query (measured on B200, 20 k / 100 k Gaussians x 6890 vertices; exhaustive kernel 67 / 245 us):
// one thread per query 124 / 85 us -- every thread walks its cells through a chain of dependent, uncoalesced loads;
// 32 lanes 25 / 95 us -- ~900 warp instructions per query for ~60 candidates per ring; 8 lanes 31 / 77 us; 16 lanes
// 24 / 82 us.  Per ring the lanes of a group take one (z, y) row of the block each -- the cells of a row are consecutive,
// hence so are their vertices in `sorted`: two loads give a row's whole range -- the rows' ranges are flattened into one
// candidate list by a scan over the group, every lane takes candidates k = lane, lane + 16, ... (all loads of a ring
// independent: two round trips per ring), and a lexicographic shuffle-min merges the lanes.  The groups of a warp run
// their own trip counts; every shuffle names its group's lanes only.
constexpr int GRID_GW = 16;                       // lanes per query
constexpr int GRID_QUERY_THREADS = 256;           // 32 queries per block

__global__ void __launch_bounds__(GRID_QUERY_THREADS)
nearest_vertex_grid_kernel(int P, const float* __restrict__ xyz, const VertexGridView g, int* __restrict__ vert_ids, float* __restrict__ dist)
{
	const int lane = threadIdx.x & 31, gl = lane & (GRID_GW - 1);
	const unsigned gmask = ((1u << GRID_GW) - 1u) << (lane & ~(GRID_GW - 1));
	const int idx = (blockIdx.x * GRID_QUERY_THREADS + threadIdx.x) / GRID_GW;
	if (idx >= P) return;                                               // whole groups leave together
	const float qx = __ldg(xyz + 3 * (size_t)idx), qy = __ldg(xyz + 3 * (size_t)idx + 1), qz = __ldg(xyz + 3 * (size_t)idx + 2);
	const float lx = g.header[0], ly = g.header[1], lz = g.header[2], h = g.header[3], inv_h = g.header[4];
	const int V = __float_as_int(g.header[5]);
	const float INF = __int_as_float(0x7f800000);
	float bd = INF;
	int bi = 0x7fffffff;
	if (!(fabsf(qx) < INF && fabsf(qy) < INF && fabsf(qz) < INF)) {     // NaN / infinite query: what the exhaustive kernel answers
		if (gl == 0) { vert_ids[idx] = 0; if (dist) dist[idx] = INF; }
		return;
	}
	const int cx = grid_coord(qx, lx, inv_h), cy = grid_coord(qy, ly, inv_h), cz = grid_coord(qz, lz, inv_h);
	for (int r = 0; r < GRID_G; r++) {
		// the shell of ring r, row by row: rows on a z / y face take their whole x range, interior rows only their two ends
		const int x0 = cx - r, x1 = cx + r, y0 = cy - r, y1 = cy + r, z0 = cz - r, z1 = cz + r;
		const int xa = max(x0, 0), xb = min(x1, GRID_G - 1);
		const int side = 2 * r + 1, nrows = side * side;
		for (int rb = 0; rb < nrows; rb += GRID_GW) {
			const int row = rb + gl;
			const int z = z0 + row / side, y = y0 + row % side;
			int s0 = 0, n0 = 0, s1 = 0, n1 = 0;
			if (row < nrows && z >= 0 && z < GRID_G && y >= 0 && y < GRID_G) {
				const int* cs = g.cell_start + (z * GRID_G + y) * GRID_G;
				if (z == z0 || z == z1 || y == y0 || y == y1) {
					s0 = __ldg(cs + xa); n0 = __ldg(cs + xb + 1) - s0;
				} else {
					if (x0 >= 0) { s0 = __ldg(cs + x0); n0 = __ldg(cs + x0 + 1) - s0; }
					if (x1 <= GRID_G - 1) { s1 = __ldg(cs + x1); n1 = __ldg(cs + x1 + 1) - s1; }
				}
			}
			int incl = n0 + n1;                                           // inclusive scan of the rows' candidate counts
#pragma unroll
			for (int o = 1; o < GRID_GW; o <<= 1) {
				const int t = __shfl_up_sync(gmask, incl, o, GRID_GW);
				if (gl >= o) incl += t;
			}
			const int total = __shfl_sync(gmask, incl, GRID_GW - 1, GRID_GW);
			for (int k0 = 0; k0 < total; k0 += GRID_GW) {
				const int k = k0 + gl;
				int lo = 0, hi = GRID_GW - 1;                             // owner = first lane whose inclusive count exceeds k
#pragma unroll
				for (int it = 0; it < (GRID_GW == 32 ? 5 : GRID_GW == 16 ? 4 : 3); it++) {
					const int mid = (lo + hi) >> 1;
					const int v = __shfl_sync(gmask, incl, mid, GRID_GW);
					if (v <= k) lo = mid + 1; else hi = mid;
				}
				const int owner = lo & (GRID_GW - 1);
				const int o_incl = __shfl_sync(gmask, incl, owner, GRID_GW), o_n0 = __shfl_sync(gmask, n0, owner, GRID_GW),
					o_n1 = __shfl_sync(gmask, n1, owner, GRID_GW);
				const int o_s0 = __shfl_sync(gmask, s0, owner, GRID_GW), o_s1 = __shfl_sync(gmask, s1, owner, GRID_GW);
				if (k < total) {
					const int within = k - (o_incl - o_n0 - o_n1);
					const int pos = within < o_n0 ? o_s0 + within : o_s1 + (within - o_n0);
					const float4 v = g.sorted[pos];
					const float dx = __fsub_rn(qx, v.x), dy = __fsub_rn(qy, v.y), dz = __fsub_rn(qz, v.z);
					const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
					nearer(bd, bi, d2, __float_as_int(v.w));
				}
			}
		}
		// the group's best so far, on every lane of the group
		float wd = bd;
		int wi = bi;
#pragma unroll
		for (int o = GRID_GW / 2; o > 0; o >>= 1) {
			const float od = __shfl_xor_sync(gmask, wd, o, GRID_GW);
			const int oi = __shfl_xor_sync(gmask, wi, o, GRID_GW);
			nearer(wd, wi, od, oi);
		}
		bd = wd; bi = wi;
		// everything inside the block [c - r, c + r]^3 has been seen.  An unseen vertex lies beyond one of the block's faces
		// that are still inside the grid; the nearest such face bounds its distance from below.
		float bound = INF;
		if (x0 > 0) bound = fminf(bound, qx - (lx + (float)x0 * h));
		if (x1 < GRID_G - 1) bound = fminf(bound, (lx + (float)(x1 + 1) * h) - qx);
		if (y0 > 0) bound = fminf(bound, qy - (ly + (float)y0 * h));
		if (y1 < GRID_G - 1) bound = fminf(bound, (ly + (float)(y1 + 1) * h) - qy);
		if (z0 > 0) bound = fminf(bound, qz - (lz + (float)z0 * h));
		if (z1 < GRID_G - 1) bound = fminf(bound, (lz + (float)(z1 + 1) * h) - qz);
		if (bound == INF) break;                                          // the block covers the grid
		const float safe = bound * 0.9999f - 1e-7f * (fabsf(qx) + fabsf(qy) + fabsf(qz) + 1.f);     // fp32 slack of cell assignment and d2
		if (safe > 0.f && bd <= safe * safe) break;
	}
	if (gl == 0) {
		vert_ids[idx] = (bi >= 0 && bi < V) ? bi : 0;
		if (dist) dist[idx] = __fsqrt_rn(bd);
	}
}

} // namespace moss
