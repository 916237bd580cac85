// K4a: per-tile segment sort (one 256-thread block per tile, deepest tiles first).
//
// Replaces the reference's global 64-bit radix sort over all instances + identifyTileRanges
// (DGR/cuda_rasterizer/rasterizer_impl.cu:302-320): binning.cuh already grouped the instances by tile, so every
// tile sorts its own (depth_bits<<32 | id) keys -- a total order, hence the result is bit-identical to the
// reference's stable sort of (tile|depth) keys emitted in ascending Gaussian order.
// Also holds the small device helpers shared by the blend kernels (blend.cuh).
#pragma once
#include "common.cuh"

namespace moss {

#ifdef MOSS_TILE_TIMING
// developer build only (tools/tile_timing.py): per-tile phase clocks, [tile][8]
__device__ long long g_tile_timing[16384 * 8];
__device__ long long g_tile_detail[16384 * 16];
__device__ __forceinline__ long long moss_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return (long long)t; }
#define MOSS_TT(slot) do { if (threadIdx.x == 0) g_tile_timing[tile * 8 + (slot)] = moss_gtime(); } while (0)
#define MOSS_TD(slot) do { if (threadIdx.x == 0) g_tile_detail[tile * 16 + (slot)] = clock64(); } while (0)
#else
#define MOSS_TT(slot) do { } while (0)
#define MOSS_TD(slot) do { } while (0)
#endif

// Shared-memory loads the compiler may not sink under a predicate.  nvcc turns `cond && smem[i] < x` (and even a
// select on a clamped index) into a divergent branch around the LDS, which exposes one shared-memory latency plus
// a BSSY/BSYNC pair per element; these latency-bound loops want the load issued unconditionally and masked after.
__device__ __forceinline__ unsigned long long lds_u64(const void* p)
{
	unsigned long long v;
	asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"((uint32_t)__cvta_generic_to_shared(p)));
	return v;
}
__device__ __forceinline__ float4 lds_f4(const void* p)
{
	float4 v;
	asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
		: "r"((uint32_t)__cvta_generic_to_shared(p)));
	return v;
}

constexpr int SORT_CAP = 6144;   // instances per tile sorted in shared memory (48 KB of keys + 12 KB of bucket bytes,
                                 // dynamic shared memory; side views of the ZJU body reach ~4400 per tile)
constexpr int NPATCH = TILE_PIX / 32;   // 8 warps = 2 (x) by 4 (y) patches of 8x4 pixels

// Ascending bitonic network over a[0..n) with virtual +inf padding up to the next power of two
// ("flip" formulation: every compare-exchange is ascending, so padding never moves).
template <typename KeyPtr>
__device__ __forceinline__ void block_bitonic_sort(KeyPtr a, int n)
{
	int npow2 = 1;
	while (npow2 < n) npow2 <<= 1;
	const int half = npow2 >> 1;
	for (int k = 2; k <= npow2; k <<= 1) {
		for (int j = k >> 1; j > 0; j >>= 1) {
			const bool flip = (j == (k >> 1));
			for (int t = threadIdx.x; t < half; t += TILE_PIX) {
				const int blk = t / j, off = t - blk * j;
				const int i = blk * 2 * j + off;
				const int p = flip ? (blk * 2 * j + 2 * j - 1 - off) : (i + j);
				if (p < n) {
					const unsigned long long x = a[i], y = a[p];
					if (x > y) { a[i] = y; a[p] = x; }
				}
			}
			__syncthreads();
		}
	}
}

// power = -0.5f*(A dx^2 + C dy^2) - B dx dy  (DGR forward.cu:340, backward.cu:507) in the rounding sequence of
// the reference's sm_100 SASS: q = fma(dx, dx*A, dy*(dy*C)); power = fma(q, -0.5, -(dy*(dx*B))).
__device__ __forceinline__ float blend_power(float dx, float dy, const float4& co)
{
	const float q = ffma(dx, fmul(dx, co.x), fmul(dy, fmul(dy, co.z)));
	return ffma(q, -0.5f, -fmul(dy, fmul(dx, co.y)));
}

// ---- per-tile sort -----------------------------------------------------------------------------------
// Sorts the tile's n (depth_bits<<32 | id) keys (all distinct: ids are) and writes the sorted id list (and, for
// parity exports, the reference's 64-bit tile|depth keys).
//   n <= 256           : bitonic network in shared memory (<= 36 compare-exchange stages);
//   256 < n <= SORT_CAP: SAMPLE SORT, O(n) work: the depth words of 255 evenly spaced keys are rank-sorted and
//                        used as splitters; every key finds its bucket by binary search, buckets are laid out
//                        by a counting sort (shared-memory atomics), and the final position of a key is its
//                        bucket's start + the number of smaller keys in that bucket (expected bucket size
//                        n/256 <= 16, observed maximum ~45).  The deepest ZJU tile (1704 keys) takes ~22k
//                        cycles; the bitonic network took ~8x that, and it sits on the launch's critical path;
//   n > SORT_CAP       : in-place bitonic network on the global segment (rare; correctness fallback).
// Ranks of the 256 splitter samples (thread t holds the depth word of sample t; ties broken by t, so the ranks are a
// permutation): samp[rank] = mine.  Every warp sorts its 32 (depth, index) pairs with a shuffle bitonic network, the
// eight sorted runs go to shared memory (`runs`, 256 x 8 bytes of scratch), and a sample's rank is its position in its
// own run plus, for each of the seven other runs, the number of smaller elements found by a 6-probe binary search:
// ~350 instructions per thread.  The all-pairs count this replaces (each thread against all 256 samples, ~1600
// instructions) was 40 % of the sort kernel's instructions at Stress size (ncu: 248 M warp instructions for 8160 tiles).
// Contains one __syncthreads(); the caller synchronises before reading samp.
__device__ __forceinline__ void rank_splitters(uint32_t mine, unsigned long long* runs, uint32_t* samp)
{
	const unsigned full = 0xffffffffu;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	unsigned long long k = ((unsigned long long)mine << 8) | (unsigned)tid;
#pragma unroll
	for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
		for (int stride = size >> 1; stride > 0; stride >>= 1) {
			const unsigned long long other = __shfl_xor_sync(full, k, stride);
			const bool keep_min = ((lane & size) == 0) == ((lane & stride) == 0);   // ascending half keeps the minimum in its lower lane
			const bool other_less = other < k;
			k = (keep_min == other_less) ? other : k;
		}
	}
	runs[warp * 32 + lane] = k;
	__syncthreads();
	int r = lane;
#pragma unroll
	for (int w = 0; w < TILE_PIX / 32; w++) {
		const unsigned long long* a = runs + w * 32;
		int lo = 0;
#pragma unroll
		for (int st = 16; st > 0; st >>= 1) lo += (lds_u64(a + lo + st - 1) < k) ? st : 0;
		lo += (lds_u64(a + lo) < k) ? 1 : 0;
		r += (w != warp) ? lo : 0;
	}
	samp[r] = (uint32_t)(k >> 8);
}

struct SortSmem {
	unsigned long long keys[SORT_CAP];   // bucketed keys (sample sort) / the keys themselves (bitonic)
	uint32_t samp[256];                  // sorted depth words of the splitters
	uint32_t start[257];
	uint32_t cur[256];
	uint32_t wsum[8];
	uint8_t bucket[SORT_CAP];            // bucket of key i (by position in the unsorted segment)
	uint8_t bslot[SORT_CAP];             // bucket of the key stored in slot p of the bucketed array
};

__device__ __forceinline__ void emit_sorted(const BinWS& b, uint32_t dst, int tile, unsigned long long k)
{
	b.point_list[dst] = (uint32_t)k;
	if (b.keys_sorted) b.keys_sorted[dst] = ((unsigned long long)tile << 32) | (k >> 32);
}

// Tiles deeper than SORT_CAP (edge-on views of a 100 k-Gaussian body reach 8-12 k instances in one tile): the same
// sample sort with the bucketed keys in GLOBAL scratch (BinWS::keys_tmp) and only the per-key bucket bytes in shared
// memory; every bucket (expected size n/256) is then ranked by one warp.  The in-place global bitonic network this
// replaces needed ~100 barrier-separated passes over L2 (0.5-0.9 ms for such a tile).
constexpr int LARGE_SORT_CAP = SORT_CAP * 8;      // one bucket byte per key in the 48 KB key area

__device__ __forceinline__ void tile_sort_large(SortSmem& sm, const BinWS& b, uint2 range, int n, int tile)
{
	const unsigned long long* gk = b.keys + range.x;
	unsigned long long* tmp = b.keys_tmp + range.x;
	uint8_t* bucket_of = reinterpret_cast<uint8_t*>(sm.keys);          // [n] bytes, n <= 8 * SORT_CAP
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	// 1. splitters (sorted depth words of 255 evenly spaced keys), ranked exactly like the shared-memory version
	const uint32_t mine = tid < 255 ? (uint32_t)(gk[(int)(((long long)(tid + 1) * n) >> 8)] >> 32) : 0xffffffffu;
	sm.cur[tid] = 0u;
	rank_splitters(mine, sm.keys, sm.samp);                            // sm.keys: scratch here, the bucket bytes from step 2 on
	__syncthreads();
	// 2. bucket of every key, bucket sizes
	for (int i = tid; i < n; i += TILE_PIX) {
		const uint32_t d = (uint32_t)(gk[i] >> 32);
		int pos = 0;
#pragma unroll
		for (int st = 128; st > 0; st >>= 1) pos += (sm.samp[pos + st - 1] < d) ? st : 0;
		bucket_of[i] = (uint8_t)pos;
		atomicAdd(&sm.cur[pos], 1u);
	}
	__syncthreads();
	// 3. exclusive scan over the 256 bucket sizes
	{
		const uint32_t c = sm.cur[tid];
		uint32_t v = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
			if (lane >= o) v += nb;
		}
		if (lane == 31) sm.wsum[warp] = v;
		__syncthreads();
		uint32_t off = 0;
#pragma unroll
		for (int w = 0; w < 8; w++) off += (w < warp) ? sm.wsum[w] : 0u;
		const uint32_t excl = off + v - c;
		__syncthreads();                       // everyone has read sm.start as scratch / sm.cur as counts
		sm.start[tid] = excl;
		sm.cur[tid] = excl;
		if (tid == 255) sm.start[256] = (uint32_t)n;
	}
	__syncthreads();
	// 4. scatter into the buckets (global scratch)
	for (int i = tid; i < n; i += TILE_PIX) tmp[atomicAdd(&sm.cur[bucket_of[i]], 1u)] = gk[i];
	__syncthreads();
	// 5. one warp per bucket: final position = bucket start + number of smaller keys in the bucket
	for (int bk = warp; bk < 256; bk += TILE_PIX / 32) {
		const uint32_t s0 = sm.start[bk], len = sm.start[bk + 1] - s0;
		for (uint32_t j = lane; j < len; j += 32) {
			const unsigned long long key = tmp[s0 + j];
			uint32_t r = 0;
			for (uint32_t q = 0; q < len; q++) r += (tmp[s0 + q] < key) ? 1u : 0u;
			emit_sorted(b, range.x + s0 + r, tile, key);
		}
	}
}

template <int SORT_ILP>
__device__ __forceinline__ void sample_sort_tail(SortSmem& sm, const BinWS& b, uint2 range, int n, int tile, const unsigned long long* gk)
{
	// Steps 2, 4 and 5 handle SORT_ILP keys per thread at a time so their global loads, binary searches and
	// bucket scans overlap (each step is a chain of dependent shared-memory reads per key).
	const int tid = threadIdx.x;
	// 2. bucket = number of splitters below the key (0..255); bucket sizes
	for (int c0 = 0; c0 < n; c0 += SORT_ILP * TILE_PIX) {
		unsigned long long key[SORT_ILP];
		int pos[SORT_ILP];
#pragma unroll
		for (int u = 0; u < SORT_ILP; u++) {
			const int i = c0 + u * TILE_PIX + tid;
			key[u] = i < n ? gk[i] : ~0ull;
			pos[u] = 0;
		}
#pragma unroll
		for (int st = 128; st > 0; st >>= 1) {
#pragma unroll
			for (int u = 0; u < SORT_ILP; u++)
				pos[u] += (sm.samp[pos[u] + st - 1] < (uint32_t)(key[u] >> 32)) ? st : 0;
		}
#pragma unroll
		for (int u = 0; u < SORT_ILP; u++) {
			const int i = c0 + u * TILE_PIX + tid;
			if (i < n) { sm.bucket[i] = (uint8_t)pos[u]; atomicAdd(&sm.cur[pos[u]], 1u); }
		}
	}
	__syncthreads();
	MOSS_TD(5);
	// 3. exclusive scan over the 256 bucket sizes
	{
		const int lane = tid & 31, warp = tid >> 5;
		const uint32_t c = sm.cur[tid];
		uint32_t v = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
			if (lane >= o) v += nb;
		}
		if (lane == 31) sm.wsum[warp] = v;
		__syncthreads();
		uint32_t off = 0;
#pragma unroll
		for (int w = 0; w < 8; w++) off += (w < warp) ? sm.wsum[w] : 0u;
		const uint32_t excl = off + v - c;
		sm.start[tid] = excl;
		sm.cur[tid] = excl;
		if (tid == 255) sm.start[256] = (uint32_t)n;
	}
	__syncthreads();
	// 4. counting-sort scatter into the buckets
	for (int c0 = 0; c0 < n; c0 += SORT_ILP * TILE_PIX) {
#pragma unroll
		for (int u = 0; u < SORT_ILP; u++) {
			const int i = c0 + u * TILE_PIX + tid;
			if (i < n) {
				const uint8_t bk = sm.bucket[i];
				const uint32_t slot = atomicAdd(&sm.cur[bk], 1u);
				sm.keys[slot] = gk[i];
				sm.bslot[slot] = bk;
			}
		}
	}
	__syncthreads();
	MOSS_TD(6);
	// 5. final position = bucket start + rank inside the bucket.  Slot-centric: consecutive lanes hold
	//    consecutive slots, i.e. mostly the same bucket, so the scan reads are near-broadcast (a key-centric
	//    scan issues random-address LDS.64 and saturates the shared-memory pipe with bank conflicts).
	for (int c0 = 0; c0 < n; c0 += SORT_ILP * TILE_PIX) {
		unsigned long long key[SORT_ILP];
		uint32_t st[SORT_ILP], len[SORT_ILP], r[SORT_ILP];
		uint32_t maxlen = 0;
#pragma unroll
		for (int u = 0; u < SORT_ILP; u++) {
			const int p = c0 + u * TILE_PIX + tid;
			key[u] = 0ull; st[u] = 0u; len[u] = 0u; r[u] = 0u;
			if (p < n) {
				key[u] = sm.keys[p];
				const int bk = sm.bslot[p];
				st[u] = sm.start[bk];
				len[u] = sm.start[bk + 1] - st[u];
			}
			maxlen = max(maxlen, len[u]);
		}
		for (uint32_t q = 0; q < maxlen; q++) {
#pragma unroll
			for (int u = 0; u < SORT_ILP; u++) {
				// branch-free: read a clamped slot, mask the result (a guarded read compiles to a divergent
				// branch per u and exposes one shared-memory latency each: ~800 cycles per q in the first version)
				const unsigned long long other = lds_u64(&sm.keys[min(st[u] + q, (uint32_t)n - 1u)]);
				r[u] += ((q < len[u]) & (other < key[u])) ? 1u : 0u;
			}
		}
#pragma unroll
		for (int u = 0; u < SORT_ILP; u++) {
			const int p = c0 + u * TILE_PIX + tid;
			if (p < n) emit_sorted(b, range.x + st[u] + r[u], tile, key[u]);
		}
	}
}

__device__ __forceinline__ void tile_sort(SortSmem& sm, const BinWS& b, uint2 range, int n, int tile)
{
	unsigned long long* gk = b.keys + range.x;
	const int tid = threadIdx.x;
	if (n <= TILE_PIX) {
		if (tid < n) sm.keys[tid] = gk[tid];
		__syncthreads();
		block_bitonic_sort(sm.keys, n);
		if (tid < n) emit_sorted(b, range.x + tid, tile, sm.keys[tid]);
		return;
	}
	if (n > SORT_CAP) {
		if (n <= LARGE_SORT_CAP && b.keys_tmp) { tile_sort_large(sm, b, range, n, tile); return; }
		__syncthreads();
		block_bitonic_sort(gk, n);
		for (int i = tid; i < n; i += TILE_PIX) emit_sorted(b, range.x + i, tile, gk[i]);
		return;
	}
	MOSS_TD(8);
	// 1. splitters: the depth words of 255 evenly spaced keys, sorted by rank (rank_splitters: warp-sorted runs merged
	//    by binary searches; a 36-stage block-wide bitonic network costs ~5x the all-pairs count this replaced).
	//    Buckets are defined on the depth word only: bucket(k) = #splitters < depth(k) is monotone in the full
	//    key, which is all the counting sort needs; the ranking inside a bucket compares full keys.
	const uint32_t mine = tid < 255 ? (uint32_t)(gk[(int)(((long long)(tid + 1) * n) >> 8)] >> 32) : 0xffffffffu;
	sm.cur[tid] = 0u;
	rank_splitters(mine, sm.keys, sm.samp);                     // sm.keys is scratch here: the bucketed keys are written only in step 4
	__syncthreads();
	MOSS_TD(4);
	// Steps 2-5 with as many keys per thread in flight as the tile needs: a chunk is SORT_ILP x 256 slots, and a 900-key
	// tile walked in chunks of 2048 spends half of its binary searches and bucket scans on slots that hold nothing.
	if (n <= 2 * TILE_PIX) sample_sort_tail<2>(sm, b, range, n, tile, gk);
	else if (n <= 4 * TILE_PIX) sample_sort_tail<4>(sm, b, range, n, tile, gk);
	else sample_sort_tail<8>(sm, b, range, n, tile, gk);
	MOSS_TD(9);
}

// One block per tile: block b sorts the (b / V)-th deepest tile of view b % V (deepest tiles of every view first).
// `tile` handed to the sort is the tile's index INSIDE its view (what the reference's 64-bit keys carry).
#ifndef MOSS_SORT_MAXREG
#define MOSS_SORT_MAXREG 0
#endif
#if MOSS_SORT_MAXREG > 0
#define MOSS_SORT_REGCAP __maxnreg__(MOSS_SORT_MAXREG)
#else
#define MOSS_SORT_REGCAP __launch_bounds__(TILE_PIX)
#endif
__global__ void MOSS_SORT_REGCAP
tile_sort_kernel(ImageWS im, BinWS b, CkptWS ck, int T, int V)
{
	extern __shared__ __align__(16) unsigned char sort_smem_raw[];
	SortSmem& ss = *reinterpret_cast<SortSmem*>(sort_smem_raw);
	__shared__ uint32_t seg_off;
	if (im.counters[1]) return;   // binning overflow: the host re-runs with a larger buffer
	const uint32_t view = blockIdx.x % (uint32_t)V, rank = blockIdx.x / (uint32_t)V;
	if (rank >= im.view_busy[view]) return;                   // only non-empty tiles have work
	const int vtile = (int)im.tile_order[(size_t)view * T + rank];   // virtual tile v*T + t
	const int tile = vtile - (int)view * T;
	const uint2 range = im.ranges[vtile];
	const int n = (int)(range.y - range.x);
	if (ck.rec && n > 0) {
		// hand the tile its backward segments and boundary records.  Blocks start deepest tiles first, so the deep
		// tiles' segments land at the front of seg_map; the exact order only affects scheduling.
		const uint32_t nseg = ((uint32_t)n + CKPT_SEG - 1) / CKPT_SEG;
		if (threadIdx.x == 0) {
			seg_off = atomicAdd(&im.counters[4], nseg);
			ck.seg_base[vtile] = atomicAdd(&im.counters[5], nseg - 1u);
		}
		__syncthreads();
		for (uint32_t i = threadIdx.x; i < nseg; i += TILE_PIX)
			if (seg_off + i < ck.max_segs) ck.seg_map[seg_off + i] = make_uint2((uint32_t)vtile, i);
	}
	MOSS_TT(0);
	if (n > 0) tile_sort(ss, b, range, n, tile);
	MOSS_TT(1);
}

} // namespace moss
