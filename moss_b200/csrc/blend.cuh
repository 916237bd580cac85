// K4b: front-to-back alpha blend (forward);  K6: reverse-order blend backward.
//
// Blend semantics: DGR/cuda_rasterizer/forward.cu:261-383 (renderCUDA) and backward.cu:399-587.
// The reference runs one 256-thread block per 16x16 tile and lets every pixel walk every instance of the tile.
// At ZJU shape only ~200 tiles are busy and the deepest ones (1000-1700 instances) set the launch time, so this
// version is organised around the critical path of ONE deep tile:
//  * a warp owns an 8x4 pixel patch; a block is NW warps (NW = 8, 4 or 2: whole, half or quarter tile).  With
//    NW < 8 the patches of a deep tile run on different SMs, each warp has a scheduler to itself (the hit loop is
//    issue-bound at two warps per scheduler) and every sub-block stops as soon as ITS pixels are saturated;
//  * PATCH CULLING: while a batch of instances is staged in shared memory, the staging thread tests its instance
//    against the block's patches (exact minimum of the conic's quadratic form over the patch rectangle versus the
//    alpha >= 1/255 threshold, conservative margin).  Each warp compacts the slots that can reach its patch into a
//    byte list and walks only those (~1/3 of the tile's instances at ZJU splat sizes).  A culled instance is one
//    every pixel of the patch would have skipped through `alpha < 1/255` (forward.cu:350), so image, final_T and
//    n_contrib stay bit-identical (tests: test_patch_culling_is_invisible, MOSS_FLAG_NO_CULL);
//  * the hit loop is branch-free: FWD_GROUP hits are fetched with one broadcast load of their slot bytes, their
//    alphas are evaluated independently (ILP), then a short predicated chain T -> T(1-alpha) -> compare runs
//    per hit.  (Guarded loads compile to divergent branches; each exposed a shared-memory latency: 4x slower.)
//  * colour+depth are staged with the conic (one float4 gather per instance) and the gathers of batch k+1 are in
//    flight while batch k is blended;
//  * backward: same culling/compaction, gradients reduced across the warp with a recursive-halving shuffle tree,
//    stored (not atomically added) into per-warp rows in shared memory, summed over the block's warps and flushed
//    with ONE global atomic per (instance, block, component) per batch; batches behind the block's last
//    contributor are never loaded.
// The forward arithmetic uses explicitly-rounded intrinsics in the reference's contraction order so the
// image is bit-identical on identical sorted lists (thresholds at power>0, alpha<1/255, T<1e-4 make a 1-ulp
// difference visible).
#pragma once
#include "common.cuh"
#include "tile_sort.cuh"

namespace moss {

// Patch p (0..7) of a tile: column p&1 (8 px wide), row p>>1 (4 px high).  Block `sub` of a tile split into
// 8/NW blocks owns patches sub*NW .. sub*NW+NW-1.
template <int NW>
__device__ __forceinline__ void pixel_of_thread(int tile_x, int tile_y, int sub, int& px, int& py)
{
	const int patch = sub * NW + (threadIdx.x >> 5), lane = threadIdx.x & 31;
	px = tile_x * TILE_X + (patch & 1) * 8 + (lane & 7);
	py = tile_y * TILE_Y + (patch >> 1) * 4 + (lane >> 3);
}

// Bit w set <=> the instance may reach alpha >= 1/255 on some pixel of the block's w-th patch.
// alpha = o*exp(-q/2) >= 1/255  <=>  q(d) = A dx^2 + 2B dx dy + C dy^2 <= 2 ln(255 o).  The minimum of the convex q
// over the patch's offset rectangle [dxl,dxh]x[dyl,dyh] is 0 if the centre is inside; otherwise it lies on an edge
// facing the centre, i.e. at (ex, clamp(-B ex/C)) or (clamp(-B ey/A), ey) with (ex,ey) the rectangle point
// nearest to 0 per axis.  Both are feasible points, so min(q1,q2) is exact up to rounding; the margin
// (0.02 absolute on q, 3e-5 relative on the positive terms) is >30x the fp32 error of either evaluation and
// of expf/logf, so the test never culls an instance the exact per-pixel test would blend.
// NaN anywhere -> comparison false -> not culled.
template <int NW>
__device__ __forceinline__ uint32_t patch_hit_mask(const float2 xy, const float4 co, float tile_px0, float tile_py0, int sub)
{
	static_assert(NW == 2 || NW == 4 || NW == 8, "a block owns whole patch rows");
	const float A = co.x, B = co.y, Cc = co.z;
	const float thr = 2.0f * __logf(255.0f * co.w) + 0.02f;     // fast log: |error| << the 0.02 margin
	const float nbc = __fdividef(-B, Cc), nba = __fdividef(-B, A);
	const float B2 = 2.0f * B;
	const float mx = xy.x - tile_px0;
	constexpr int ROWS = NW / 2;
	const float my = xy.y - tile_py0 - 4.0f * (float)(sub * ROWS);
	uint32_t m = 0;
#pragma unroll
	for (int r = 0; r < ROWS; r++) {
		const float dyh = my - 4.0f * r, dyl = dyh - 3.0f;
		const float ey = fminf(fmaxf(0.0f, dyl), dyh);
		const float cey2 = Cc * ey * ey;
		const float tx = nba * ey;
#pragma unroll
		for (int c = 0; c < 2; c++) {
			const float dxh = mx - 8.0f * c, dxl = dxh - 7.0f;
			const float ex = fminf(fmaxf(0.0f, dxl), dxh);
			const float dy1 = fminf(fmaxf(nbc * ex, dyl), dyh);
			const float dx2 = fminf(fmaxf(tx, dxl), dxh);
			const float m1 = A * ex * ex + Cc * dy1 * dy1;
			const float m2 = A * dx2 * dx2 + cey2;
			const float q1 = m1 + B2 * ex * dy1 - 3e-5f * m1;
			const float q2 = m2 + B2 * dx2 * ey - 3e-5f * m2;
			if (!(fminf(q1, q2) > thr)) m |= 1u << (r * 2 + c);
		}
	}
	return m;
}

// instances staged in shared memory per round: 128, or one per thread for the 256-thread whole-tile block
template <int NW> struct BlendCfg { static constexpr int NT = 32 * NW, BATCH = NT > 128 ? NT : 128, PER = BATCH / NT; };
constexpr int FWD_GROUP = 8;       // hits whose alpha is evaluated together before the serial transmittance chain

template <int NW>
struct BlendSmem {
	static constexpr int BATCH = BlendCfg<NW>::BATCH;
	float2 xy[BATCH];
	float4 co[BATCH];
	float4 feat[BATCH];
	uint8_t mask[BATCH];
	// per warp: slots of the staged batch that can reach the warp's patch, in list order (+ pad for the last group)
	__align__(8) uint8_t hits[NW][BATCH + FWD_GROUP];
};

// Per-warp compaction of the batch's hit slots into hits[]; returns their number.  One ballot per 32 slots.
template <int BATCH>
__device__ __forceinline__ int compact_hits(const uint8_t* mask, uint8_t* hits, int cnt, int warp, int lane)
{
	int nh = 0;
#pragma unroll
	for (int sb = 0; sb < BATCH; sb += 32) {
		const bool mine = (sb + lane < cnt) && ((mask[sb + lane] >> warp) & 1u);
		const uint32_t bits = __ballot_sync(0xffffffffu, mine);
		if (mine) hits[nh + __popc(bits & ((1u << lane) - 1u))] = (uint8_t)(sb + lane);
		nh += __popc(bits);
	}
	__syncwarp();
	return nh;
}

// Register caps (A/B on B200, tools/ab_libs.sh, profiles/r02_h_ab_regcaps.txt): the compiler's own choice is 96-98
// registers forward and 128 backward.  80 forward costs nothing (no spills) and buys a resident block: Stress blend
// forward 412 -> 369 us.  Backward: 96 is spill-free; 80 spills 16-64 bytes but still wins where the blocks are large
// (4 / 8 warps: 23-view batch 634 -> 595 us, Stress 1004 -> 950 us) and loses on the 2-warp blocks of a single ZJU view
// (54.5 -> 56.7 us), which therefore stay at 96.  64 forward spills and loses 11 % at ZJU.  -DMOSS_BLEND_*_MAXREG overrides.
#ifndef MOSS_BLEND_FWD_MAXREG
#define MOSS_BLEND_FWD_MAXREG 80
#endif
#ifndef MOSS_BLEND_BWD_MAXREG
#define MOSS_BLEND_BWD_MAXREG (NW == 2 ? 96 : 80)
#endif
#define MOSS_FWD_REGCAP __maxnreg__(MOSS_BLEND_FWD_MAXREG)
#define MOSS_BWD_REGCAP __maxnreg__(MOSS_BLEND_BWD_MAXREG)

template <bool CULL, int NW>
__global__ void MOSS_FWD_REGCAP
blend_forward_kernel(const __grid_constant__ ViewBatch vb, GeomWS g, ImageWS im, BinWS b, CkptWS ck,
	float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_alpha)
{
	constexpr int NT = BlendCfg<NW>::NT;          // threads per block
	constexpr int BLEND_BATCH = BlendCfg<NW>::BATCH;
	constexpr int PER = BlendCfg<NW>::PER;        // instances staged per thread per batch
	constexpr int SPLIT = NPATCH / NW;            // blocks per tile
	__shared__ BlendSmem<NW> bs;

	if (im.counters[1]) return;   // binning overflow: the host re-runs with a larger buffer
	const int W = vb.W, H = vb.H, grid_x = vb.grid_x;
	// unit u = blockIdx.x / SPLIT: view u % V, that view's (u / V)-th deepest tile (deepest tiles of all views first)
	const uint32_t unit = blockIdx.x / SPLIT;
	const int view = (int)(unit % (uint32_t)vb.V);
	const int vtile = (int)im.tile_order[(size_t)view * vb.T + unit / (uint32_t)vb.V];   // virtual tile v*T + t
	const int tile = vtile - view * vb.T;
	const int sub = blockIdx.x % SPLIT;
	const int tile_y = tile / grid_x, tile_x = tile - tile_y * grid_x;
	const uint2 range = im.ranges[vtile];
	const int n = (int)(range.y - range.x);
	const unsigned full = 0xffffffffu;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

	int px, py;
	pixel_of_thread<NW>(tile_x, tile_y, sub, px, py);
	const bool inside = px < W && py < H;
	const float pxf = (float)px, pyf = (float)py;
	const float tile_px0 = (float)(tile_x * TILE_X), tile_py0 = (float)(tile_y * TILE_Y);
	bool done = !inside;
	float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, Dp = 0.f;
	uint32_t last_contributor = 0;
#ifdef MOSS_TILE_TIMING
	if (sub == 0) MOSS_TT(2);
#endif

	// software pipeline: the gathers of batch k+1 are in flight while batch k is blended
	float2 nxy[PER];
	float4 nco[PER], nfeat[PER];
#pragma unroll
	for (int u = 0; u < PER; u++) {
		const int i = u * NT + (int)threadIdx.x;
		nxy[u] = make_float2(0.f, 0.f); nco[u] = make_float4(0.f, 0.f, 0.f, 0.f); nfeat[u] = nco[u];
		if (i < n) {
			const uint32_t id = b.point_list[range.x + i];
			nxy[u] = g.means2D[id]; nco[u] = g.conic_opacity[id]; nfeat[u] = g.feat[id];
		}
	}

	// boundary records of this tile (nullptr: no checkpointing); pixel slot = patch * 32 + lane, whatever NW is
	float* ck_rec = ck.rec ? ck.rec + (size_t)ck.seg_base[vtile] * CKPT_REC_FLOATS + (sub * NW + warp) * 32 + lane : nullptr;

	for (int base = 0; base < n; base += BLEND_BATCH) {
		if (__syncthreads_count(done) == NT) break;
		if (ck_rec && base > 0 && (base % CKPT_SEG) == 0) {
			// state in front of position `base`; a pixel that is already done never has it read back
			// (its last contributor lies in front of the boundary)
			float* r = ck_rec + (size_t)(base / CKPT_SEG - 1) * CKPT_REC_FLOATS;
			r[0] = T; r[TILE_PIX] = C0; r[2 * TILE_PIX] = C1; r[3 * TILE_PIX] = C2; r[4 * TILE_PIX] = Dp;
		}
#pragma unroll
		for (int u = 0; u < PER; u++) {
			const int s = u * NT + (int)threadIdx.x;
			uint32_t hit = 0;
			if (base + s < n) {
				bs.xy[s] = nxy[u];
				bs.co[s] = nco[u];
				bs.feat[s] = nfeat[u];
				hit = CULL ? patch_hit_mask<NW>(nxy[u], nco[u], tile_px0, tile_py0, sub) : 0xffu;
			}
			bs.mask[s] = (uint8_t)hit;
		}
		__syncthreads();
#pragma unroll
		for (int u = 0; u < PER; u++) {
			const int i = base + BLEND_BATCH + u * NT + (int)threadIdx.x;
			if (i < n) {
				const uint32_t id = b.point_list[range.x + i];
				nxy[u] = g.means2D[id]; nco[u] = g.conic_opacity[id]; nfeat[u] = g.feat[id];
			}
		}
		if (__all_sync(full, done)) continue;          // this warp's patch is finished (still helps staging)
		const int cnt = min(BLEND_BATCH, n - base);
		const int nh = compact_hits<BLEND_BATCH>(bs.mask, bs.hits[warp], cnt, warp, lane);

		// Software pipeline over groups of FWD_GROUP hits: the alphas of group g+1 (independent work: loads, power,
		// exp) are evaluated in the same straight-line block as the serial transmittance chain of group g, so the
		// chain's dependent latency is covered by the next group's arithmetic.
		int jj[FWD_GROUP];
		float al[FWD_GROUP];
		bool ok[FWD_GROUP];
		auto eval_group = [&](int h0, int (&oj)[FWD_GROUP], float (&oa)[FWD_GROUP], bool (&ook)[FWD_GROUP]) {
			const unsigned long long packed = lds_u64(&bs.hits[warp][h0]);
#pragma unroll
			for (int k = 0; k < FWD_GROUP; k++) {
				// pad bytes are stale; masked to a valid slot
				const int j = (int)((packed >> (8 * k)) & (unsigned long long)(BLEND_BATCH - 1));
				oj[k] = j;
				const float2 xy = bs.xy[j];
				const float4 co = bs.co[j];
				const float dx = fsub(xy.x, pxf), dy = fsub(xy.y, pyf);
				const float power = blend_power(dx, dy, co);
				const float alpha = fminf(fmul(co.w, expf(power)), 0.99f);
				oa[k] = alpha;
				ook[k] = (h0 + k < nh) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
			}
		};
		if (nh > 0) eval_group(0, jj, al, ok);
		for (int h = 0; h < nh; h += FWD_GROUP) {
			int njj[FWD_GROUP];
			float nal[FWD_GROUP];
			bool nok[FWD_GROUP];
			// (reads past the list end stay inside hits[warp][] thanks to its FWD_GROUP pad; results are masked)
			eval_group(min(h + FWD_GROUP, BLEND_BATCH), njj, nal, nok);
			// serial part, branch-free: the feature rows are fetched up front (independent LDS) and every
			// update is predicated, so a hit costs one short dependent chain T -> test_T -> compare
			float4 f[FWD_GROUP];
#pragma unroll
			for (int k = 0; k < FWD_GROUP; k++) f[k] = lds_f4(&bs.feat[jj[k]]);
#pragma unroll
			for (int k = 0; k < FWD_GROUP; k++) {
				const float alpha = al[k];
				const float test_T = fmul(T, fsub(1.0f, alpha));
				const bool live = ok[k] && !done;
				const bool stop = live && (test_T < 0.0001f);
				const bool upd = live && !stop;
				done = done || stop;
				const float c0 = ffma(T, fmul(alpha, f[k].x), C0);
				const float c1 = ffma(T, fmul(alpha, f[k].y), C1);
				const float c2 = ffma(T, fmul(alpha, f[k].z), C2);
				const float wn = ffma(T, alpha, weight);
				const float dn = ffma(T, fmul(alpha, f[k].w), Dp);
				C0 = upd ? c0 : C0; C1 = upd ? c1 : C1; C2 = upd ? c2 : C2;
				weight = upd ? wn : weight; Dp = upd ? dn : Dp;
				T = upd ? test_T : T;
				last_contributor = upd ? (uint32_t)(base + jj[k] + 1) : last_contributor;
			}
#pragma unroll
			for (int k = 0; k < FWD_GROUP; k++) { jj[k] = njj[k]; al[k] = nal[k]; ok[k] = nok[k]; }
			if (__all_sync(full, done)) break;
		}
	}
#ifdef MOSS_TILE_TIMING
	if (sub == 0) MOSS_TT(3);
#endif

	if (inside) {
		const int pix = py * W + px;
		const size_t HW = (size_t)H * W;
		const size_t vpix = (size_t)view * HW + pix;          // pixel of the batch: outputs are [V,3,H,W] / [V,H,W]
		const float* __restrict__ bg = vb.cam[view].bg;
		im.final_T[vpix] = T;
		im.n_contrib[vpix] = last_contributor;
		float* oc = out_color + (size_t)view * 3 * HW;
		oc[pix] = ffma(T, bg[0], C0);
		oc[HW + pix] = ffma(T, bg[1], C1);
		oc[2 * HW + pix] = ffma(T, bg[2], C2);
		out_alpha[vpix] = weight;   // sum of alpha*T, DGR forward.cu:380
		out_depth[vpix] = Dp;
		if (ck.fin) ck.fin[vpix] = make_float4(C0, C1, C2, Dp);
	}
}

// Sum v[0..7] over the 32 lanes with recursive halving: afterwards the 4 lanes sharing (lane>>2) hold the
// total of value (lane>>2).  9 shuffles instead of 40.
__device__ __forceinline__ float warp_reduce8(const float (&v)[8])
{
	const unsigned full = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	float w[4], u[2];
	const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
#pragma unroll
	for (int i = 0; i < 4; i++) {
		const float send = b4 ? v[i] : v[i + 4];
		const float keep = b4 ? v[i + 4] : v[i];
		w[i] = keep + __shfl_xor_sync(full, send, 16);
	}
#pragma unroll
	for (int i = 0; i < 2; i++) {
		const float send = b3 ? w[i] : w[i + 2];
		const float keep = b3 ? w[i + 2] : w[i];
		u[i] = keep + __shfl_xor_sync(full, send, 8);
	}
	const float send = b2 ? u[0] : u[1];
	const float keep = b2 ? u[1] : u[0];
	float s = keep + __shfl_xor_sync(full, send, 4);
	s += __shfl_xor_sync(full, s, 2);
	s += __shfl_xor_sync(full, s, 1);
	return s;
}
// Sum a and b over the warp: lanes 0-15 end with the total of a, lanes 16-31 with the total of b (5 shuffles).
__device__ __forceinline__ float warp_reduce2(float a, float b)
{
	const unsigned full = 0xffffffffu;
	const bool hi = threadIdx.x & 16;
	float s = (hi ? b : a) + __shfl_xor_sync(full, hi ? a : b, 16);
#pragma unroll
	for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(full, s, o);
	return s;
}

// Transpose-reduce: sum v[0..31] over the 32 lanes with recursive halving; afterwards lane l holds the total of
// value l.  31 shuffles for 32 values (a butterfly per value would need 160), 5 dependent stages.
__device__ __forceinline__ float warp_reduce32(const float (&v)[32])
{
	const unsigned full = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
	float a[16], c[8], d[4], e[2];
#pragma unroll
	for (int i = 0; i < 16; i++) {
		const float send = b4 ? v[i] : v[i + 16];
		const float keep = b4 ? v[i + 16] : v[i];
		a[i] = keep + __shfl_xor_sync(full, send, 16);
	}
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const float send = b3 ? a[i] : a[i + 8];
		const float keep = b3 ? a[i + 8] : a[i];
		c[i] = keep + __shfl_xor_sync(full, send, 8);
	}
#pragma unroll
	for (int i = 0; i < 4; i++) {
		const float send = b2 ? c[i] : c[i + 4];
		const float keep = b2 ? c[i + 4] : c[i];
		d[i] = keep + __shfl_xor_sync(full, send, 4);
	}
#pragma unroll
	for (int i = 0; i < 2; i++) {
		const float send = b1 ? d[i] : d[i + 2];
		const float keep = b1 ? d[i + 2] : d[i];
		e[i] = keep + __shfl_xor_sync(full, send, 2);
	}
	const float send = b0 ? e[0] : e[1];
	const float keep = b0 ? e[1] : e[0];
	return keep + __shfl_xor_sync(full, send, 1);
}
// Sum v[0..3] over the warp: the 8 lanes sharing (lane>>3) end with the total of value (lane>>3).  6 shuffles.
__device__ __forceinline__ float warp_reduce4(const float (&v)[4])
{
	const unsigned full = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const bool b4 = lane & 16, b3 = lane & 8;
	float a[2];
#pragma unroll
	for (int i = 0; i < 2; i++) {
		const float send = b4 ? v[i] : v[i + 2];
		const float keep = b4 ? v[i + 2] : v[i];
		a[i] = keep + __shfl_xor_sync(full, send, 16);
	}
	const float send = b3 ? a[0] : a[1];
	const float keep = b3 ? a[1] : a[0];
	float s = keep + __shfl_xor_sync(full, send, 8);
	s += __shfl_xor_sync(full, s, 4);
	s += __shfl_xor_sync(full, s, 2);
	s += __shfl_xor_sync(full, s, 1);
	return s;
}

constexpr int NGRAD = 9;          // dcolor[3], dmean2D.xy, dconic.{x,y,w}, dopacity
constexpr int BWD_BATCH = 128;    // instances staged per round of the backward
constexpr int BWD_GROUP = 4;      // hits processed together: their independent halves interleave and their 4 x 8
                                  // gradient terms go through ONE 32-value transpose-reduce (5 dependent shuffle stages per 4 hits)

template <int NW>
struct BwdSmem {
	uint32_t id[BWD_BATCH];
	float2 xy[BWD_BATCH];
	float4 co[BWD_BATCH];
	float4 feat[BWD_BATCH];
	uint8_t mask[BWD_BATCH];
	uint8_t written[NW][BWD_BATCH];                   // warp w stored a gradient row for this slot (plain byte stores)
	__align__(8) uint8_t hits[NW][BWD_BATCH + 8];
	uint32_t warp_last[NW];
	float grad[NW * BWD_BATCH * NGRAD];               // [warp][slot][component]: each (warp, slot) is written at
	                                                  // most once per batch -> plain stores, no shared atomics
};

template <int NW>
__global__ void MOSS_BWD_REGCAP
blend_backward_kernel(const __grid_constant__ ViewBatch vb, GeomWS g, ImageWS im, BinWS b, CkptWS ck,
	const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepths, const float* __restrict__ dL_dalphas,
	int deterministic)
{
	constexpr int NT = 32 * NW;
	constexpr int SPLIT = NPATCH / NW;
	constexpr int PER = (BWD_BATCH + NT - 1) / NT;
	__shared__ BwdSmem<NW> sm;

	// Unit of work: a whole tile, or -- with a checkpoint workspace -- one CKPT_SEG-long segment [seg_lo, seg_hi) of
	// its instance list, so that the segments of a deep tile run as independent blocks.
	const int W = vb.W, H = vb.H, grid_x = vb.grid_x;
	int vtile, seg = 0;              // virtual tile v*T + t
	if (ck.rec) {
		if (blockIdx.x / SPLIT >= im.counters[4]) return;
		const uint2 ts = ck.seg_map[blockIdx.x / SPLIT];
		vtile = (int)ts.x; seg = (int)ts.y;
	} else {
		const uint32_t unit = blockIdx.x / SPLIT;           // deepest tiles of every view first (binning.cuh)
		const uint32_t v_ = unit % (uint32_t)vb.V;
		vtile = (int)im.tile_order[(size_t)v_ * vb.T + unit / (uint32_t)vb.V];
	}
	const int view = vtile / vb.T;
	const int tile = vtile - view * vb.T;
	const int sub = blockIdx.x % SPLIT;
	const int tile_y = tile / grid_x, tile_x = tile - tile_y * grid_x;
	const uint2 range = im.ranges[vtile];
	const int n = (int)(range.y - range.x);
	if (n == 0) return;
	const float* __restrict__ bg = vb.cam[view].bg;
	const uint32_t seg_lo = ck.rec ? (uint32_t)seg * CKPT_SEG : 0u;
	const uint32_t seg_hi = ck.rec ? min((uint32_t)n, seg_lo + CKPT_SEG) : (uint32_t)n;

	int px, py;
	pixel_of_thread<NW>(tile_x, tile_y, sub, px, py);
	const bool inside = px < W && py < H;
	const int pix = py * W + px;
	const float pxf = (float)px, pyf = (float)py;
	const float tile_px0 = (float)(tile_x * TILE_X), tile_py0 = (float)(tile_y * TILE_Y);
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned full = 0xffffffffu;

	const size_t HW = (size_t)H * W;
	const size_t vpix = (size_t)view * HW + pix;
	const float T_final = inside ? im.final_T[vpix] : 0.f;
	float T = T_final;
	const uint32_t n_contrib = inside ? im.n_contrib[vpix] : 0u;
	// positions [seg_lo, last_contributor) of this unit can receive gradient from this pixel
	const uint32_t last_contributor = n_contrib > seg_lo ? min(n_contrib, seg_hi) : seg_lo;
	float dpix[3] = {0.f, 0.f, 0.f}, ddepth = 0.f, dalpha = 0.f;
	if (inside) {
		const float* dp = dL_dpix + (size_t)view * 3 * HW;
		dpix[0] = dp[pix]; dpix[1] = dp[HW + pix]; dpix[2] = dp[2 * HW + pix];
		ddepth = dL_ddepths ? dL_ddepths[vpix] : 0.f;      // NULL = no gradient reaches this output
		dalpha = dL_dalphas ? dL_dalphas[vpix] : 0.f;
	}
	const float bg_dot_dpixel = bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2];

	// nothing behind the furthest "last contributor" of the block (of the warp) can receive gradient
	const uint32_t warp_last = __reduce_max_sync(full, last_contributor);
	if (lane == 0) sm.warp_last[warp] = warp_last;
	__syncthreads();
	uint32_t block_last = 0;
#pragma unroll
	for (int w = 0; w < NW; w++) block_last = max(block_last, sm.warp_last[w]);
	if (block_last <= seg_lo) return;
#ifdef MOSS_TILE_TIMING
	if (sub == 0) MOSS_TT(4);
#endif

	// State of the back-to-front walk.  The reference keeps, per channel, (accum_rec, last_alpha, last_color) and forms
	// arc = last_alpha * last_color + (1 - last_alpha) * accum_rec at the NEXT contributing instance (DGR backward.cu:
	// 531-551).  Here the walk carries that combination itself -- accum_rec[] below IS the arc the next contributor will
	// use -- updated as alpha * c + (1 - alpha) * arc: the same operands and the same expression, one value per channel
	// instead of two.  An instance that does not contribute runs the same update with alpha = 0 (0 * c + 1 * arc = arc,
	// T / (1 - 0) = T, exactly), so the recurrences need no selects; its gradient terms vanish because they all carry
	// the masked alpha or the masked Gaussian weight as a factor.  (Per hit: 20 selects -> 2, 5 registers of state less.)
	// Assumes finite colours and depths: 0 * c of a skipped instance with a NaN / Inf colour reaches the pixel's running
	// sums here, where a select kept it out (such a Gaussian already puts NaN into every pixel it does contribute to).
	float accum_rec[3] = {0.f, 0.f, 0.f};
	float accum_depth_rec = 0.f, accum_alpha_rec = 0.f;
	if (ck.rec && n_contrib > seg_hi) {
		// the pixel kept blending behind this segment: start from the forward's state in front of position seg_hi.
		// With P = that prefix state and F the final sums, the reference's running "what lies behind" values are
		// (F - P) / T_P  (DGR backward.cu:531-551 unrolled over the instances behind the boundary).
		const float* r = ck.rec + ((size_t)ck.seg_base[vtile] + seg) * CKPT_REC_FLOATS + (sub * NW + warp) * 32 + lane;
		const float Tp = r[0];
		const float4 F = ck.fin[vpix];
		const float inv = __fdividef(1.0f, Tp);
		accum_rec[0] = (F.x - r[TILE_PIX]) * inv; accum_rec[1] = (F.y - r[2 * TILE_PIX]) * inv;
		accum_rec[2] = (F.z - r[3 * TILE_PIX]) * inv;
		accum_depth_rec = (F.w - r[4 * TILE_PIX]) * inv;
		accum_alpha_rec = (Tp - T_final) * inv;
		T = Tp;
	}
	const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
	const int n_walk = (int)(block_last - seg_lo);      // positions block_last-1 ... seg_lo, back to front

	// batch slot s holds forward position (0-based) block_last-1-(base+s): the walk is back to front.
	// Software pipeline as in the forward: the gathers of the next batch are issued before this one is processed.
	uint32_t nid[PER];
	float2 nxy[PER];
	float4 nco[PER], nfeat[PER];
#pragma unroll
	for (int u = 0; u < PER; u++) {
		const int i = u * NT + (int)threadIdx.x;
		nid[u] = 0u; nxy[u] = make_float2(0.f, 0.f); nco[u] = make_float4(0.f, 0.f, 0.f, 0.f); nfeat[u] = nco[u];
		if (u * NT + (int)threadIdx.x < BWD_BATCH && i < n_walk) {
			nid[u] = b.point_list[range.x + block_last - 1 - i];
			nxy[u] = g.means2D[nid[u]]; nco[u] = g.conic_opacity[nid[u]]; nfeat[u] = g.feat[nid[u]];
		}
	}

	for (int base = 0; base < n_walk; base += BWD_BATCH) {
		__syncthreads();
#pragma unroll
		for (int u = 0; u < PER; u++) {
			const int s = u * NT + (int)threadIdx.x;
			if (s < BWD_BATCH) {
				uint32_t hit = 0;
				if (base + s < n_walk) {
					sm.id[s] = nid[u]; sm.xy[s] = nxy[u]; sm.co[s] = nco[u]; sm.feat[s] = nfeat[u];
					hit = patch_hit_mask<NW>(nxy[u], nco[u], tile_px0, tile_py0, sub);
				}
				sm.mask[s] = (uint8_t)hit;
#pragma unroll
				for (int w = 0; w < NW; w++) sm.written[w][s] = 0;
			}
		}
		__syncthreads();
#pragma unroll
		for (int u = 0; u < PER; u++) {
			const int s = u * NT + (int)threadIdx.x;
			const int i = base + BWD_BATCH + s;
			if (s < BWD_BATCH && i < n_walk) {
				nid[u] = b.point_list[range.x + block_last - 1 - i];
				nxy[u] = g.means2D[nid[u]]; nco[u] = g.conic_opacity[nid[u]]; nfeat[u] = g.feat[nid[u]];
			}
		}

		// compact the slots that can reach this warp's patch AND lie in front of the warp's last contributor
		const int cnt = min(BWD_BATCH, n_walk - base);
		int nh = 0;
#pragma unroll
		for (int sb = 0; sb < BWD_BATCH; sb += 32) {
			const int fpos = (int)block_last - 1 - (base + sb + lane);
			const bool mine = (sb + lane < cnt) && ((sm.mask[sb + lane] >> warp) & 1u) && fpos < (int)warp_last;
			const uint32_t bits = __ballot_sync(full, mine);
			if (mine) sm.hits[warp][nh + __popc(bits & ((1u << lane) - 1u))] = (uint8_t)(sb + lane);
			nh += __popc(bits);
		}
		__syncwarp();

		for (int h = 0; h < nh; h += BWD_GROUP) {
			// ---- independent part of both hits: geometry, alpha, validity ------------------------------
			int jj[BWD_GROUP];
			float dx[BWD_GROUP], dy[BWD_GROUP], Gs[BWD_GROUP], al[BWD_GROUP], ra[BWD_GROUP];
			float4 cos_[BWD_GROUP], fs[BWD_GROUP];
			bool valid[BWD_GROUP];
#pragma unroll
			for (int k = 0; k < BWD_GROUP; k++) {
				const int j = (int)sm.hits[warp][min(h + k, nh - 1)];
				jj[k] = j;
				const float2 xy = sm.xy[j];
				const float4 co = sm.co[j];
				cos_[k] = co;
				fs[k] = lds_f4(&sm.feat[j]);
				dx[k] = xy.x - pxf; dy[k] = xy.y - pyf;
				const float power = blend_power(dx[k], dy[k], co);
				Gs[k] = expf(power);
				const float a_raw = fminf(fmul(co.w, Gs[k]), 0.99f);
				const uint32_t contributor = block_last - 1u - (uint32_t)(base + j);
				valid[k] = (h + k < nh) && contributor < last_contributor && !(power > 0.0f) && !(a_raw < 1.0f / 255.0f);
				// masked: a skipped instance has alpha 0 and weight 0 (exp(power) may be inf for power > 0: never multiplied in)
				al[k] = valid[k] ? a_raw : 0.f;
				Gs[k] = valid[k] ? Gs[k] : 0.f;
				ra[k] = __fdividef(1.0f, 1.0f - al[k]);
			}
			bool any_valid = false;
#pragma unroll
			for (int k = 0; k < BWD_GROUP; k++) any_valid = any_valid || valid[k];
			if (!__any_sync(full, any_valid)) continue;

			// ---- recurrences (serial over the hits, predicated) and gradient terms ------------------------
			float v[BWD_GROUP * 8], v_opa[BWD_GROUP];
#pragma unroll
			for (int k = 0; k < BWD_GROUP; k++) {
				const float alpha = al[k];                        // 0 for an instance that does not contribute
				const float4 f = fs[k];
				const float4 co = cos_[k];
				const float Tn = T * ra[k];                       // T / (1 - alpha), DGR backward.cu:520
				const float oma = 1.f - alpha;
				const float dchannel_dcolor = alpha * Tn;
				const float c[3] = {f.x, f.y, f.z};
				float dL_dopa = 0.0f;
#pragma unroll
				for (int ch = 0; ch < 3; ch++) {
					dL_dopa += (c[ch] - accum_rec[ch]) * dpix[ch];
					v[k * 8 + ch] = dchannel_dcolor * dpix[ch];
					accum_rec[ch] = alpha * c[ch] + oma * accum_rec[ch];
				}
				dL_dopa += (f.w - accum_depth_rec) * ddepth;
				dL_dopa += (1.f - accum_alpha_rec) * dalpha;
				dL_dopa *= Tn;
				dL_dopa += (-T_final * ra[k]) * bg_dot_dpixel;
				accum_depth_rec = alpha * f.w + oma * accum_depth_rec;
				accum_alpha_rec = alpha + oma * accum_alpha_rec;
				T = Tn;

				const float dL_dG = co.w * dL_dopa;
				const float gdx = Gs[k] * dx[k], gdy = Gs[k] * dy[k];   // Gs is 0 for a skipped instance
				const float dG_ddelx = -gdx * co.x - gdy * co.y;
				const float dG_ddely = -gdy * co.z - gdx * co.y;
				v[k * 8 + 3] = dL_dG * dG_ddelx * ddelx_dx;
				v[k * 8 + 4] = dL_dG * dG_ddely * ddely_dy;
				v[k * 8 + 5] = -0.5f * gdx * dx[k] * dL_dG;
				v[k * 8 + 6] = -0.5f * gdx * dy[k] * dL_dG;
				v[k * 8 + 7] = -0.5f * gdy * dy[k] * dL_dG;
				v_opa[k] = Gs[k] * dL_dopa;
			}
			// ---- warp reduction, one row per (warp, slot): lane l ends with component l&7 of hit l>>3 ------------
			static_assert(BWD_GROUP == 4, "the transpose-reduce below is laid out for 4 hits x 8 components");
			const float s32 = warp_reduce32(v);
			const float so = warp_reduce4(v_opa);
			const int kk = lane >> 3;
			const int jsel = kk == 0 ? jj[0] : (kk == 1 ? jj[1] : (kk == 2 ? jj[2] : jj[3]));
			if (h + kk < nh) {
				float* row = sm.grad + (warp * BWD_BATCH + jsel) * NGRAD;
				row[lane & 7] = s32;
				if ((lane & 7) == 0) { row[8] = so; sm.written[warp][jsel] = 1; }
			}
		}
		__syncthreads();

		// flush the batch: the rows of the warps that hit a slot are summed in warp order and added to the Gaussian's
		// accumulators with VECTOR atomics (red.global.add.v4.f32): a thread takes one slot and either
		// (colour rgb, opacity) or (mean2D xy, conic xy) + conic w -- 3 atomic instructions per (instance, block)
		// instead of 9
		for (int t = threadIdx.x; t < 2 * BWD_BATCH; t += NT) {
			const int j = t & (BWD_BATCH - 1);
			const bool second = t >= BWD_BATCH;
			float r[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
			bool any = false;
#pragma unroll
			for (int w = 0; w < NW; w++) {
				if (sm.written[w][j]) {
					any = true;
					const float* row = sm.grad + (w * BWD_BATCH + j) * NGRAD;
					if (!second) { r[0] += row[0]; r[1] += row[1]; r[2] += row[2]; r[3] += row[8]; }
					else { r[0] += row[3]; r[1] += row[4]; r[2] += row[5]; r[3] += row[6]; r[4] += row[7]; }
				}
			}
			if (any && !deterministic) {
				float4* acc = g.grad_acc + GRAD_ACC_VEC * (size_t)sm.id[j];
				if (!second) atomicAdd(acc, make_float4(r[0], r[1], r[2], r[3]));
				else { atomicAdd(acc + 1, make_float4(r[0], r[1], r[2], r[3])); atomicAdd(&acc[2].x, r[4]); }
			} else if (any) {
				// order-independent accumulation: the block's partial sums (themselves computed in a fixed order)
				// are added as 64-bit fixed-point integers
				const float S = det_scale_from_bits(im.counters[6], false);
				const int first = second ? 4 : 0, cnt = second ? 5 : 4;
#pragma unroll
				for (int i = 0; i < 5; i++)
					if (i < cnt) atomicAdd(det_slot(g, (size_t)sm.id[j], first + i), (unsigned long long)__float2ll_rn(r[i] * S));
			}
		}
	}
#ifdef MOSS_TILE_TIMING
	if (sub == 0) { MOSS_TT(5); if (threadIdx.x == 0) { g_tile_timing[tile * 8 + 6] = (long long)block_last; g_tile_timing[tile * 8 + 7] = (long long)n; } }
#endif
}

// max |upstream gradient| over dL_dpix [3N], dL_ddepth [N], dL_dalpha [N] -> *out (float bits, caller-zeroed).
// max is order-independent, so the fixed-point scale derived from it is reproducible.
__global__ void __launch_bounds__(256)
upstream_maxabs_kernel(int N, const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth,
	const float* __restrict__ dL_dalpha, uint32_t* __restrict__ out)
{
	// N = pixels of the whole batch (V*H*W): dL_dpix holds 3N values in any layout, max is layout-blind
	float m = 0.f;
	for (int i = blockIdx.x * 256 + threadIdx.x; i < N; i += gridDim.x * 256) {
		m = fmaxf(m, fmaxf(fmaxf(fabsf(dL_dpix[i]), fabsf(dL_dpix[N + i])), fabsf(dL_dpix[2 * N + i])));
		if (dL_ddepth) m = fmaxf(m, fabsf(dL_ddepth[i]));
		if (dL_dalpha) m = fmaxf(m, fabsf(dL_dalpha[i]));
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
	if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));   // non-negative floats order like their bits
}

} // namespace moss
