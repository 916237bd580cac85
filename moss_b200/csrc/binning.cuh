// K2: per-tile exclusive scan -> tile ranges (+ instance count);  K3: scatter (depth|id) keys into tile segments.
//
// Replaces the reference's per-Gaussian InclusiveSum + blocking D2H of num_rendered + duplicateWithKeys +
// global 64-bit CUB radix sort + memset + identifyTileRanges (DGR/cuda_rasterizer/rasterizer_impl.cu:70-138,
// 279-320).  Because K1 already counted the instances of every tile, the tile ranges are a scan over T tiles
// (one block) and every instance can be scattered straight into its tile's segment; the remaining order
// inside a segment -- (depth bits, Gaussian id), which is what a STABLE sort of (tile|depth) keys emitted in
// ascending Gaussian order produces -- is a total order, so any per-segment sort of the 64-bit
// (depth_bits<<32 | id) key reproduces the reference's sorted list bit for bit (K4, blend.cuh).
// HBM traffic per instance: 8 B write here + 8 B read + 4 B write in K4, versus ~150 B for the 6-pass
// global radix sort.
#pragma once
#include "common.cuh"

namespace moss {

constexpr int SCAN_THREADS = 1024;

// Size class of a tile for the launch order: quarter-octave buckets of its instance count (0 = empty).
constexpr int ORDER_BUCKETS = 132;
__device__ __forceinline__ int order_bucket(uint32_t c)
{
	if (c == 0) return 0;
	const int bits = 32 - __clz(c);                       // 1..32
	const int sub = bits >= 3 ? (int)((c >> (bits - 3)) & 3u) : 0;
	return bits * 4 + sub;                                // 4..131
}

// One block PER VIEW (grid = V).  Block v owns the tiles [v*T, (v+1)*T) of its view:
// ranges[t'] = [first,last) into the global sorted list (or (0,0) for an empty tile, as the reference's memset leaves
// it), tile_count[t'] <- the scatter cursors, view_busy[v] = its non-empty tiles, view_base[v] = its first instance.
// The views are chained with a one-shot look-back: block v publishes (its instance total + 1) in view_sync[v] as soon
// as it knows it and adds up the totals of the views before it (V <= 32 blocks, all resident at once).  The last view's
// block writes counters = {R, R > capacity, capacity, ...}; host_R (mapped pinned host memory, may be null) also
// receives R so the host can read it without a blocking copy.
// tile_order[v*T + k] = the k-th tile (global index) of view v in descending size class (a counting sort over
// ORDER_BUCKETS log-scale classes), empty tiles last.  Blend/sort block b works on view b % V, rank b / V: the hardware
// hands blocks to SMs in index order, so the deepest tiles of every view start first and the shallow ones fill in behind
// them (LPT scheduling); with the natural order several deep tiles of the body's centre column land on one SM and the
// launch waits for that SM while the others idle (ncu: sm__cycles_active.avg 37% of elapsed).
__global__ void __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(int T, int V, uint32_t* __restrict__ tile_count_all, uint32_t* __restrict__ tile_total_all, uint2* __restrict__ ranges_all,
	uint32_t* __restrict__ counters, uint32_t capacity, volatile int* host_R, uint32_t* __restrict__ tile_order_all,
	uint32_t* __restrict__ view_busy, uint32_t* __restrict__ view_base, uint32_t* __restrict__ view_sync)
{
	__shared__ uint32_t warp_sums[32];
	__shared__ uint32_t carry_s;
	__shared__ uint32_t bucket_n[ORDER_BUCKETS];
	const int view = blockIdx.x;
	const uint32_t tile0 = (uint32_t)view * (uint32_t)T;
	uint32_t* tile_count = tile_count_all + (size_t)tile0 * TILE_SUB;
	uint32_t* tile_total = tile_total_all + tile0;
	uint2* ranges = ranges_all + tile0;
	uint32_t* tile_order = tile_order_all + tile0;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if (threadIdx.x < ORDER_BUCKETS) bucket_n[threadIdx.x] = 0;
	__syncthreads();
	// per-tile totals; the sub-counters become the scatter cursors (exclusive prefix inside the tile's segment);
	// size-class histogram; the view's instance total
	uint32_t c_first = 0u;     // total of tile threadIdx.x, kept in a register (images of up to 1024 tiles never re-read)
	uint32_t my_sum = 0u;
	for (int t = threadIdx.x; t < T; t += SCAN_THREADS) {
		uint4* sub = reinterpret_cast<uint4*>(tile_count + (size_t)t * TILE_SUB);
		static_assert(TILE_SUB == 8, "two uint4 per tile");
		const uint4 a = sub[0], b = sub[1];
		uint4 pa, pb;
		pa.x = 0u; pa.y = a.x; pa.z = pa.y + a.y; pa.w = pa.z + a.z;
		pb.x = pa.w + a.w; pb.y = pb.x + b.x; pb.z = pb.y + b.y; pb.w = pb.z + b.z;
		const uint32_t c = pb.w + b.w;
		sub[0] = pa; sub[1] = pb;
		tile_total[t] = c;
		if (t == (int)threadIdx.x) c_first = c;
		my_sum += c;
		atomicAdd(&bucket_n[order_bucket(c)], 1u);
	}
	my_sum = __reduce_add_sync(0xffffffffu, my_sum);
	if (lane == 0) warp_sums[warp] = my_sum;
	__syncthreads();
	if (warp == 0) {
		const uint32_t view_total = __reduce_add_sync(0xffffffffu, warp_sums[lane]);
		if (lane == 0) {
			__threadfence();
			atomicExch(&view_sync[view], view_total + 1u);       // 0 = not known yet
		}
		// descending exclusive offsets: class 131 first ... class 0 (empty) last
		uint32_t run = 0;
		for (int base = ORDER_BUCKETS - 1; base >= 0; base -= 32) {
			const int bkt = base - lane;
			const uint32_t c = bkt >= 0 ? bucket_n[bkt] : 0u;
			uint32_t v = c;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
				if (lane >= o) v += nb;
			}
			if (bkt >= 0) bucket_n[bkt] = run + v - c;
			run += __shfl_sync(0xffffffffu, v, 31);
		}
		// look-back: instances of the views before this one (lane u waits for view u)
		uint32_t before = 0u;
		if (lane < view) {
			const volatile uint32_t* w = view_sync + lane;
			uint32_t x;
			while ((x = *w) == 0u) { }
			before = x - 1u;
		}
		before = __reduce_add_sync(0xffffffffu, before);
		if (lane == 0) {
			carry_s = before;
			view_base[view] = before;
			if (view == V - 1) {
				const uint32_t R = before + view_total;
				view_base[V] = R;
				counters[0] = R;
				counters[1] = (R > capacity) ? 1u : 0u;
				counters[2] = capacity;
				counters[4] = 0u;       // backward segments / checkpoint records: handed out by tile_sort_kernel
				counters[5] = 0u;
				if (host_R) { *host_R = (int)R; __threadfence_system(); }
			}
		}
	}
	__syncthreads();
	for (int t = threadIdx.x; t < T; t += SCAN_THREADS)
		tile_order[atomicAdd(&bucket_n[order_bucket(t == (int)threadIdx.x ? c_first : tile_total[t])], 1u)] = tile0 + (uint32_t)t;
	__syncthreads();
	// classes 1..3 never occur (count 1 maps to class 4), so the offset of class 1 is the number of non-empty tiles
	if (threadIdx.x == 0) {
		view_busy[view] = bucket_n[1];
		if (view == 0) counters[3] = (V == 1) ? bucket_n[1] : 0u;     // per-view counts live in view_busy
	}
	for (int base = 0; base < T; base += SCAN_THREADS) {
		const int t = base + threadIdx.x;
		const uint32_t c = (t < T) ? (base == 0 ? c_first : tile_total[t]) : 0u;
		uint32_t v = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
			if (lane >= o) v += n;
		}
		if (lane == 31) warp_sums[warp] = v;
		__syncthreads();
		if (warp == 0) {
			uint32_t w = warp_sums[lane];
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint32_t n = __shfl_up_sync(0xffffffffu, w, o);
				if (lane >= o) w += n;
			}
			warp_sums[lane] = w;
		}
		__syncthreads();
		const uint32_t carry = carry_s;
		const uint32_t incl = v + (warp > 0 ? warp_sums[warp - 1] : 0u) + carry;
		if (t < T) {
			ranges[t] = c ? make_uint2(incl - c, incl) : make_uint2(0u, 0u);
		}
		__syncthreads();
		if (threadIdx.x == SCAN_THREADS - 1) carry_s = incl;
		__syncthreads();
	}
}

// K3.  A warp flattens the tile lists of its 32 Gaussians into one index space and every lane takes one
// (Gaussian, tile) instance per round: the cursor atomics of one large splat (dozens of tiles) are issued by many
// lanes at once instead of one lane waiting for each returned slot in turn -- that serial chain of L2 round trips
// of the LARGEST splat used to set this launch's duration (22 us at ZJU shape).
// The exclusive start of tile t (ranges[t].x) is only meaningful for non-empty tiles, which is all the scatter asks for.
constexpr int SCATTER_THREADS = 128;

// grid = (ceil(P / SCATTER_THREADS), V)
__global__ void __launch_bounds__(SCATTER_THREADS)
scatter_keys_kernel(int P, int grid_x, int T, const int* __restrict__ radii, GeomWS g,
	const uint2* __restrict__ ranges, uint32_t* __restrict__ cursor, const uint32_t* __restrict__ counters,
	unsigned long long* __restrict__ keys)
{
	const unsigned full = 0xffffffffu;
	const int view = blockIdx.y;
	const int idx = blockIdx.x * SCATTER_THREADS + threadIdx.x;
	const int lane = threadIdx.x & 31;
	if (counters[1]) return;          // binning buffer too small: host re-runs with a larger one (uniform exit)
	const size_t vid = (size_t)view * P + idx;
	const uint32_t tile0 = (uint32_t)view * (uint32_t)T;
	const bool vis = idx < P && radii[vid] > 0;
	ushort4 r = make_ushort4(0, 0, 0, 0);
	uint32_t depth_bits = 0;
	if (vis) { r = g.rect[vid]; depth_bits = __float_as_uint(g.depths[vid]); }
	const uint32_t w = (uint32_t)r.z - (uint32_t)r.x;
	const uint32_t cnt = vis ? w * ((uint32_t)r.w - (uint32_t)r.y) : 0u;
	uint32_t incl = cnt;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t nb = __shfl_up_sync(full, incl, o);
		if (lane >= o) incl += nb;
	}
	const uint32_t total = __shfl_sync(full, incl, 31);
	const uint32_t origin = (uint32_t)r.x | ((uint32_t)r.y << 16);
	const uint32_t vid_lane0 = (uint32_t)(vid - lane);        // virtual ids fit 32 bits (checked by the host: V*P < 2^32)
	for (uint32_t base = 0; base < total; base += 32) {
		const uint32_t k = base + lane;
		// owner = first lane whose inclusive count exceeds k (branch-free binary search over the lanes)
		int lo = 0;
#pragma unroll
		for (int step = 16; step > 0; step >>= 1) {
			const uint32_t v = __shfl_sync(full, incl, lo + step - 1);
			if (v <= k) lo += step;
		}
		const int owner = min(lo, 31);
		const uint32_t o_incl = __shfl_sync(full, incl, owner);
		const uint32_t o_cnt = __shfl_sync(full, cnt, owner);
		const uint32_t o_w = __shfl_sync(full, w, owner);
		const uint32_t o_origin = __shfl_sync(full, origin, owner);
		const uint32_t o_depth = __shfl_sync(full, depth_bits, owner);
		if (k < total) {
			const uint32_t local = k - (o_incl - o_cnt);
			const uint32_t ty = local / o_w, tx = local - ty * o_w;
			const uint32_t t = tile0 + ((o_origin >> 16) + ty) * (uint32_t)grid_x + (o_origin & 0xffffu) + tx;
			// the cursor of the sub-counter this warp counted into (preprocess: same warp <-> same 32 Gaussians)
			const uint32_t sub = ((uint32_t)idx >> 5) & (TILE_SUB - 1);
			const uint32_t slot = ranges[t].x + atomicAdd(&cursor[(size_t)t * TILE_SUB + sub], 1u);
			// low word: the VIRTUAL Gaussian id (v*P + i): inside one view it orders like i, and the blend kernels
			// gather the view's state rows with it directly
			keys[slot] = ((unsigned long long)o_depth << 32) | (vid_lane0 + (uint32_t)owner);
		}
	}
}

} // namespace moss
