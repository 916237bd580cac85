// K4: per-tile segment sort + front-to-back alpha blend (forward);  K6: reverse-order blend backward.
//
// Blend semantics: DGR/cuda_rasterizer/forward.cu:261-383 (renderCUDA) and backward.cu:399-587.
// One 256-thread block per 16x16 tile; a warp owns an 8x4 pixel patch.  New relative to the reference:
//  * the tile's (depth|id) segment is sorted inside the block (shared memory bitonic network; in-place global
//    fallback for oversized segments) right before blending, replacing the global radix sort;
//  * colour+depth are staged in shared memory with the conic (one float4 gather per instance) instead of
//    per-pair global reads;
//  * PATCH CULLING: while a batch of 256 instances is staged, the loading thread tests its instance against the
//    tile's eight 8x4 patches (exact minimum of the conic's quadratic form over the patch rectangle versus the
//    alpha >= 1/255 threshold, with a conservative margin) and stores an 8-bit hit mask.  Each warp then walks
//    only the instances that can reach its patch (ballot -> bit scan), which cuts the serial per-pixel loop
//    ~3x at ZJU splat sizes.  A culled instance is one every pixel of the patch would have skipped through
//    `alpha < 1/255` (forward.cu:350), so the image stays bit-identical.  The masks are written next to the
//    sorted list and re-used by the backward;
//  * the alpha of up to four hits is evaluated independently before the serial transmittance chain (ILP:
//    the kernel is latency-bound, ~3 warps per scheduler at ZJU shape);
//  * backward: per-hit gradients are reduced across the warp with a recursive-halving shuffle tree, summed
//    across the tile's warps in shared memory, and flushed with ONE global atomic per (instance, tile,
//    component) after each batch; batches behind the tile's last contributor are never loaded.
// The forward arithmetic uses explicitly-rounded intrinsics in the reference's contraction order so the
// image is bit-identical on identical sorted lists (thresholds at power>0, alpha<1/255, T<1e-4 make a 1-ulp
// difference visible).
#pragma once
#include "common.cuh"

namespace moss {

constexpr int SORT_CAP = 4096;   // instances per tile sorted in shared memory (32 KB)
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

__device__ __forceinline__ void pixel_of_thread(int tile_x, int tile_y, int& px, int& py)
{
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	px = tile_x * TILE_X + (warp & 1) * 8 + (lane & 7);
	py = tile_y * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
}

// Bit w set <=> the instance may reach alpha >= 1/255 on some pixel of patch w (= warp w's 8x4 pixels).
// alpha = o*exp(-q/2) >= 1/255  <=>  q(d) = A dx^2 + 2B dx dy + C dy^2 <= 2 ln(255 o).  The minimum of the convex q
// over the patch's offset rectangle [dxl,dxh]x[dyl,dyh] is 0 if the centre is inside; otherwise it lies on an edge
// facing the centre, i.e. at (ex, clamp(-B ex/C)) or (clamp(-B ey/A), ey) with (ex,ey) the rectangle point
// nearest to 0 per axis.  Both are feasible points, so min(q1,q2) is exact up to rounding; the margin
// (0.02 absolute on q, 3e-5 relative on the positive terms) is >30x the fp32 error of either evaluation and
// of expf/logf, so the test never culls an instance the exact per-pixel test would blend.
// NaN anywhere -> comparison false -> not culled.
__device__ __forceinline__ uint32_t patch_hit_mask(const float2 xy, const float4 co, float tile_px0, float tile_py0)
{
	const float A = co.x, B = co.y, Cc = co.z;
	const float thr = 2.0f * logf(255.0f * co.w) + 0.02f;
	const float nbc = __fdividef(-B, Cc), nba = __fdividef(-B, A);
	const float B2 = 2.0f * B;
	const float mx = xy.x - tile_px0, my = xy.y - tile_py0;
	uint32_t m = 0;
#pragma unroll
	for (int r = 0; r < 4; r++) {
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

template <bool CULL>
__global__ void __launch_bounds__(TILE_PIX)
blend_forward_kernel(int W, int H, int grid_x, const float* __restrict__ bg, GeomWS g, ImageWS im, BinWS b,
	float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_alpha)
{
	__shared__ unsigned long long s_keys[SORT_CAP];
	__shared__ float2 s_xy[TILE_PIX];
	__shared__ float4 s_co[TILE_PIX];
	__shared__ float4 s_feat[TILE_PIX];
	__shared__ uint8_t s_mask[TILE_PIX];

	if (im.counters[1]) return;   // binning overflow: the host re-runs with a larger buffer
	const int tile = blockIdx.x;
	const int tile_y = tile / grid_x, tile_x = tile - tile_y * grid_x;
	const uint2 range = im.ranges[tile];
	const int n = (int)(range.y - range.x);
	const bool in_smem = n <= SORT_CAP;
	const unsigned full = 0xffffffffu;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

	// ---- sort this tile's segment by (depth bits, gaussian id) --------------------------------------
	if (n > 0) {
		unsigned long long* gk = b.keys + range.x;
		if (in_smem) {
			for (int i = threadIdx.x; i < n; i += TILE_PIX) s_keys[i] = gk[i];
			__syncthreads();
			block_bitonic_sort(s_keys, n);
			for (int i = threadIdx.x; i < n; i += TILE_PIX) {
				const unsigned long long k = s_keys[i];
				b.point_list[range.x + i] = (uint32_t)k;
				if (b.keys_sorted) b.keys_sorted[range.x + i] = ((unsigned long long)tile << 32) | (k >> 32);
			}
		} else {
			__syncthreads();
			block_bitonic_sort(gk, n);
			for (int i = threadIdx.x; i < n; i += TILE_PIX) {
				const unsigned long long k = gk[i];
				b.point_list[range.x + i] = (uint32_t)k;
				if (b.keys_sorted) b.keys_sorted[range.x + i] = ((unsigned long long)tile << 32) | (k >> 32);
			}
		}
	}

	// ---- blend ---------------------------------------------------------------------------------------
	int px, py;
	pixel_of_thread(tile_x, tile_y, px, py);
	const bool inside = px < W && py < H;
	const float pxf = (float)px, pyf = (float)py;
	const float tile_px0 = (float)(tile_x * TILE_X), tile_py0 = (float)(tile_y * TILE_Y);
	bool done = !inside;
	float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, Dp = 0.f;
	uint32_t last_contributor = 0;

	for (int base = 0; base < n; base += TILE_PIX) {
		if (__syncthreads_count(done) == TILE_PIX) break;
		const int i = base + threadIdx.x;
		uint32_t hit = 0;
		if (i < n) {
			const uint32_t id = in_smem ? (uint32_t)s_keys[i] : (uint32_t)b.keys[range.x + i];
			const float2 xy = g.means2D[id];
			const float4 co = g.conic_opacity[id];
			s_xy[threadIdx.x] = xy;
			s_co[threadIdx.x] = co;
			s_feat[threadIdx.x] = g.feat[id];
			hit = CULL ? patch_hit_mask(xy, co, tile_px0, tile_py0) : 0xffu;
			b.hit_mask[range.x + i] = (uint8_t)hit;
		}
		s_mask[threadIdx.x] = (uint8_t)hit;
		__syncthreads();
		if (__all_sync(full, done)) continue;          // this warp's patch is finished (still helps staging)
		const int cnt = min(TILE_PIX, n - base);
		for (int sb = 0; sb < cnt; sb += 32) {
			uint32_t bits = __ballot_sync(full, (s_mask[sb + lane] >> warp) & 1u);
			while (bits) {
				// up to four hits: independent alpha evaluations first, then the serial transmittance chain
				int jj[4];
				float al[4];
				bool ok[4];
#pragma unroll
				for (int k = 0; k < 4; k++) {
					const bool has = bits != 0u;
					const int j = sb + (has ? (__ffs(bits) - 1) : 0);
					bits &= bits - 1u;
					jj[k] = j;
					const float2 xy = s_xy[j];
					const float4 co = s_co[j];
					const float dx = fsub(xy.x, pxf), dy = fsub(xy.y, pyf);
					const float power = blend_power(dx, dy, co);
					const float alpha = fminf(fmul(co.w, expf(power)), 0.99f);
					al[k] = alpha;
					ok[k] = has && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
				}
#pragma unroll
				for (int k = 0; k < 4; k++) {
					if (ok[k] && !done) {
						const float alpha = al[k];
						const float test_T = fmul(T, fsub(1.0f, alpha));
						if (test_T < 0.0001f) { done = true; }
						else {
							const float4 f = s_feat[jj[k]];
							C0 = ffma(T, fmul(alpha, f.x), C0);
							C1 = ffma(T, fmul(alpha, f.y), C1);
							C2 = ffma(T, fmul(alpha, f.z), C2);
							weight = ffma(T, alpha, weight);
							Dp = ffma(T, fmul(alpha, f.w), Dp);
							T = test_T;
							last_contributor = (uint32_t)(base + jj[k] + 1);
						}
					}
				}
				if (__all_sync(full, done)) { bits = 0u; sb = cnt; }
			}
		}
	}

	if (inside) {
		const int pix = py * W + px;
		im.final_T[pix] = T;
		im.n_contrib[pix] = last_contributor;
		const int HW = H * W;
		out_color[pix] = ffma(T, bg[0], C0);
		out_color[HW + pix] = ffma(T, bg[1], C1);
		out_color[2 * HW + pix] = ffma(T, bg[2], C2);
		out_alpha[pix] = weight;   // sum of alpha*T, DGR forward.cu:380
		out_depth[pix] = Dp;
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
__device__ __forceinline__ float warp_reduce1(float v)
{
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

constexpr int NGRAD = 9;          // dcolor[3], dmean2D.xy, dconic.{x,y,w}, dopacity
constexpr int BWD_BATCH = 128;    // instances staged per round of the backward

__global__ void __launch_bounds__(TILE_PIX)
blend_backward_kernel(int W, int H, int grid_x, const float* __restrict__ bg, GeomWS g, ImageWS im, BinWS b,
	const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepths, const float* __restrict__ dL_dalphas,
	float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
	float* __restrict__ dL_dcolor)
{
	__shared__ uint32_t s_id[BWD_BATCH];
	__shared__ float2 s_xy[BWD_BATCH];
	__shared__ float4 s_co[BWD_BATCH];
	__shared__ float4 s_feat[BWD_BATCH];
	__shared__ uint8_t s_mask[BWD_BATCH];
	__shared__ uint32_t s_written[BWD_BATCH];                 // bit w: warp w stored a gradient row for this slot
	__shared__ float s_grad[NPATCH * BWD_BATCH * NGRAD];      // [warp][slot][component]: each (warp, slot) is written
	                                                          // at most once per batch -> plain stores, no atomics
	__shared__ uint32_t s_warp_last[NPATCH];

	const int tile = blockIdx.x;
	const int tile_y = tile / grid_x, tile_x = tile - tile_y * grid_x;
	const uint2 range = im.ranges[tile];
	const int n = (int)(range.y - range.x);
	if (n == 0) return;

	int px, py;
	pixel_of_thread(tile_x, tile_y, px, py);
	const bool inside = px < W && py < H;
	const int pix = py * W + px;
	const float pxf = (float)px, pyf = (float)py;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned full = 0xffffffffu;

	const float T_final = inside ? im.final_T[pix] : 0.f;
	float T = T_final;
	const uint32_t last_contributor = inside ? im.n_contrib[pix] : 0u;
	float dpix[3] = {0.f, 0.f, 0.f}, ddepth = 0.f, dalpha = 0.f;
	if (inside) {
		const int HW = H * W;
		dpix[0] = dL_dpix[pix]; dpix[1] = dL_dpix[HW + pix]; dpix[2] = dL_dpix[2 * HW + pix];
		ddepth = dL_ddepths[pix];
		dalpha = dL_dalphas[pix];
	}
	const float bg_dot_dpixel = bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2];

	// nothing behind the furthest "last contributor" of the tile (of the warp) can receive gradient
	const uint32_t warp_last = __reduce_max_sync(full, last_contributor);
	if (lane == 0) s_warp_last[warp] = warp_last;
	__syncthreads();
	uint32_t tile_last = 0;
#pragma unroll
	for (int w = 0; w < NPATCH; w++) tile_last = max(tile_last, s_warp_last[w]);
	if (tile_last == 0) return;

	float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
	float accum_depth_rec = 0.f, last_depth = 0.f, accum_alpha_rec = 0.f, last_alpha = 0.f;
	const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

	// batch slot j holds forward position (0-based) tile_last-1-(base+j): the walk is back to front
	for (int base = 0; base < (int)tile_last; base += BWD_BATCH) {
		__syncthreads();
		const int i = base + threadIdx.x;
		uint32_t hit = 0;
		if (threadIdx.x < BWD_BATCH && i < (int)tile_last) {
			const uint32_t pos = range.x + tile_last - 1 - i;
			const uint32_t id = b.point_list[pos];
			hit = b.hit_mask[pos];
			s_id[threadIdx.x] = id;
			s_xy[threadIdx.x] = g.means2D[id];
			s_co[threadIdx.x] = g.conic_opacity[id];
			s_feat[threadIdx.x] = g.feat[id];
		}
		if (threadIdx.x < BWD_BATCH) { s_mask[threadIdx.x] = (uint8_t)hit; s_written[threadIdx.x] = 0u; }
		__syncthreads();

		const int cnt = min(BWD_BATCH, (int)tile_last - base);
		for (int sb = 0; sb < cnt; sb += 32) {
			// forward position of slot sb+lane; instances at/behind the warp's last contributor get no gradient
			const int fpos = (int)tile_last - 1 - (base + sb + lane);
			uint32_t bits = __ballot_sync(full, ((s_mask[sb + lane] >> warp) & 1u) && fpos < (int)warp_last);
			while (bits) {
				const int j = sb + __ffs(bits) - 1;
				bits &= bits - 1u;
				const uint32_t contributor = tile_last - 1 - (uint32_t)(base + j);
				bool valid = contributor < last_contributor;
				const float2 xy = s_xy[j];
				const float4 co = s_co[j];
				const float dx = xy.x - pxf, dy = xy.y - pyf;
				const float power = blend_power(dx, dy, co);
				valid = valid && !(power > 0.0f);
				const float G = expf(power);
				const float alpha = fminf(fmul(co.w, G), 0.99f);
				valid = valid && !(alpha < 1.0f / 255.0f);
				if (!__any_sync(full, valid)) continue;

				float v[8], v_opa = 0.f;
#pragma unroll
				for (int k = 0; k < 8; k++) v[k] = 0.f;
				if (valid) {
					const float4 f = s_feat[j];
					T = T / (1.f - alpha);
					const float dchannel_dcolor = alpha * T;
					float dL_dopa = 0.0f;
					const float c[3] = {f.x, f.y, f.z};
#pragma unroll
					for (int ch = 0; ch < 3; ch++) {
						accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
						last_color[ch] = c[ch];
						dL_dopa += (c[ch] - accum_rec[ch]) * dpix[ch];
						v[ch] = dchannel_dcolor * dpix[ch];
					}
					accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
					last_depth = f.w;
					dL_dopa += (f.w - accum_depth_rec) * ddepth;
					accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
					dL_dopa += (1.f - accum_alpha_rec) * dalpha;
					dL_dopa *= T;
					last_alpha = alpha;
					dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;

					const float dL_dG = co.w * dL_dopa;
					const float gdx = G * dx, gdy = G * dy;
					const float dG_ddelx = -gdx * co.x - gdy * co.y;
					const float dG_ddely = -gdy * co.z - gdx * co.y;
					v[3] = dL_dG * dG_ddelx * ddelx_dx;
					v[4] = dL_dG * dG_ddely * ddely_dy;
					v[5] = -0.5f * gdx * dx * dL_dG;
					v[6] = -0.5f * gdx * dy * dL_dG;
					v[7] = -0.5f * gdy * dy * dL_dG;
					v_opa = G * dL_dopa;
				}
				const float s8 = warp_reduce8(v);
				const float s1 = warp_reduce1(v_opa);
				float* row = s_grad + (warp * BWD_BATCH + j) * NGRAD;
				if ((lane & 3) == 0) row[lane >> 2] = s8;
				else if (lane == 1) { row[8] = s1; atomicOr(&s_written[j], 1u << warp); }
			}
		}
		__syncthreads();

		// flush the batch: the rows of the warps that hit a slot are summed in warp order, then ONE global atomic
		// per (instance, component); thread t takes slot t%128 and components 0-4 (t<128) or 5-8
		{
			const int j = threadIdx.x & (BWD_BATCH - 1);
			const uint32_t wmask = s_written[j];
			if (wmask) {
				const int k0 = (threadIdx.x < BWD_BATCH) ? 0 : 5;
				float r[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
				for (int w = 0; w < NPATCH; w++) {
					if (wmask & (1u << w)) {
						const float* row = s_grad + (w * BWD_BATCH + j) * NGRAD + k0;
#pragma unroll
						for (int k = 0; k < 4; k++) r[k] += row[k];
						if (k0 == 0) r[4] += row[4];
					}
				}
				const size_t id = s_id[j];
				if (k0 == 0) {
					atomicAdd(dL_dcolor + 3 * id + 0, r[0]);
					atomicAdd(dL_dcolor + 3 * id + 1, r[1]);
					atomicAdd(dL_dcolor + 3 * id + 2, r[2]);
					atomicAdd(dL_dmean2D + 3 * id + 0, r[3]);
					atomicAdd(dL_dmean2D + 3 * id + 1, r[4]);
				} else {
					atomicAdd(dL_dconic + 4 * id + 0, r[0]);
					atomicAdd(dL_dconic + 4 * id + 1, r[1]);
					atomicAdd(dL_dconic + 4 * id + 3, r[2]);
					atomicAdd(dL_dopacity + id, r[3]);
				}
			}
		}
	}
}

} // namespace moss
