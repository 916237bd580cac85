// K4: per-tile segment sort + front-to-back alpha blend (forward);  K6: reverse-order blend backward.
//
// Blend semantics: DGR/cuda_rasterizer/forward.cu:261-383 (renderCUDA) and backward.cu:399-587.
// One 256-thread block per 16x16 tile.  New relative to the reference:
//  * the tile's (depth|id) segment is sorted inside the block (shared memory bitonic network; in-place global
//    fallback for oversized segments) right before blending, replacing the global radix sort;
//  * colour+depth are staged in shared memory with the conic (one float4 gather per instance) instead of
//    per-pair global reads;
//  * a warp covers an 8x4 pixel patch (not 16x2) so whole warps skip Gaussians that miss their patch;
//  * backward: gradients are reduced across the warp with a recursive-halving shuffle tree (14 shuffles for
//    9 values instead of 45) and only then hit global atomics (8 atomics per warp per Gaussian, issued by
//    8 different lanes), and batches behind the tile's last contributor are never loaded.
// The forward arithmetic uses explicitly-rounded intrinsics in the reference's contraction order so the
// image is bit-identical on identical sorted lists (thresholds at power>0, alpha<1/255, T<1e-4 make a 1-ulp
// difference visible).
#pragma once
#include "common.cuh"

namespace moss {

constexpr int SORT_CAP = 4096;   // instances per tile sorted in shared memory (32 KB)

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

__global__ void __launch_bounds__(TILE_PIX)
blend_forward_kernel(int W, int H, int grid_x, const float* __restrict__ bg, GeomWS g, ImageWS im, BinWS b,
	float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_alpha)
{
	__shared__ unsigned long long s_keys[SORT_CAP];
	__shared__ float2 s_xy[TILE_PIX];
	__shared__ float4 s_co[TILE_PIX];
	__shared__ float4 s_feat[TILE_PIX];

	if (im.counters[1]) return;   // binning overflow: the host re-runs with a larger buffer
	const int tile = blockIdx.x;
	const int tile_y = tile / grid_x, tile_x = tile - tile_y * grid_x;
	const uint2 range = im.ranges[tile];
	const int n = (int)(range.y - range.x);
	const bool in_smem = n <= SORT_CAP;

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
	bool done = !inside;
	float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, Dp = 0.f;
	uint32_t contributor = 0, last_contributor = 0;

	for (int base = 0; base < n; base += TILE_PIX) {
		if (__syncthreads_count(done) == TILE_PIX) break;
		const int i = base + threadIdx.x;
		if (i < n) {
			const uint32_t id = in_smem ? (uint32_t)s_keys[i] : (uint32_t)b.keys[range.x + i];
			s_xy[threadIdx.x] = g.means2D[id];
			s_co[threadIdx.x] = g.conic_opacity[id];
			s_feat[threadIdx.x] = g.feat[id];
		}
		__syncthreads();
		const int cnt = min(TILE_PIX, n - base);
		for (int j = 0; !done && j < cnt; j++) {
			contributor++;
			const float2 xy = s_xy[j];
			const float4 co = s_co[j];
			const float dx = fsub(xy.x, pxf), dy = fsub(xy.y, pyf);
			const float power = blend_power(dx, dy, co);
			if (power > 0.0f) continue;
			const float alpha = fminf(fmul(co.w, expf(power)), 0.99f);
			if (alpha < 1.0f / 255.0f) continue;
			const float test_T = fmul(T, fsub(1.0f, alpha));
			if (test_T < 0.0001f) { done = true; continue; }
			const float4 f = s_feat[j];
			C0 = ffma(T, fmul(alpha, f.x), C0);
			C1 = ffma(T, fmul(alpha, f.y), C1);
			C2 = ffma(T, fmul(alpha, f.z), C2);
			weight = ffma(T, alpha, weight);
			Dp = ffma(T, fmul(alpha, f.w), Dp);
			T = test_T;
			last_contributor = contributor;
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

__global__ void __launch_bounds__(TILE_PIX)
blend_backward_kernel(int W, int H, int grid_x, const float* __restrict__ bg, GeomWS g, ImageWS im, BinWS b,
	const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepths, const float* __restrict__ dL_dalphas,
	float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
	float* __restrict__ dL_dcolor)
{
	__shared__ uint32_t s_id[TILE_PIX];
	__shared__ float2 s_xy[TILE_PIX];
	__shared__ float4 s_co[TILE_PIX];
	__shared__ float4 s_feat[TILE_PIX];
	__shared__ uint32_t s_warp_last[TILE_PIX / 32];

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
	const uint32_t warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
	if (lane == 0) s_warp_last[warp] = warp_last;
	__syncthreads();
	uint32_t tile_last = 0;
#pragma unroll
	for (int w = 0; w < TILE_PIX / 32; w++) tile_last = max(tile_last, s_warp_last[w]);
	if (tile_last == 0) return;

	float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
	float accum_depth_rec = 0.f, last_depth = 0.f, accum_alpha_rec = 0.f, last_alpha = 0.f;
	const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

	// forward position (1-based) of the instance handled next, walking back to front
	uint32_t contributor = tile_last;
	for (int base = 0; base < (int)tile_last; base += TILE_PIX) {
		__syncthreads();
		const int i = base + threadIdx.x;   // distance behind position tile_last-1
		if (i < (int)tile_last) {
			const uint32_t id = b.point_list[range.x + tile_last - 1 - i];
			s_id[threadIdx.x] = id;
			s_xy[threadIdx.x] = g.means2D[id];
			s_co[threadIdx.x] = g.conic_opacity[id];
			s_feat[threadIdx.x] = g.feat[id];
		}
		__syncthreads();
		const int cnt = min(TILE_PIX, (int)tile_last - base);
		for (int j = 0; j < cnt; j++) {
			contributor--;
			if (contributor >= warp_last) continue;   // warp-uniform
			bool valid = contributor < last_contributor;
			const float2 xy = s_xy[j];
			const float4 co = s_co[j];
			const float dx = xy.x - pxf, dy = xy.y - pyf;
			const float power = blend_power(dx, dy, co);
			valid = valid && !(power > 0.0f);
			const float G = expf(power);
			const float alpha = fminf(fmul(co.w, G), 0.99f);
			valid = valid && !(alpha < 1.0f / 255.0f);
			if (!__any_sync(0xffffffffu, valid)) continue;

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
			const uint32_t id = s_id[j];
			if ((lane & 3) == 0) {
				const int k = lane >> 2;
				float* dst;
				if (k < 3) dst = dL_dcolor + 3 * (size_t)id + k;
				else if (k < 5) dst = dL_dmean2D + 3 * (size_t)id + (k - 3);
				else dst = dL_dconic + 4 * (size_t)id + (k == 7 ? 3 : k - 5);
				atomicAdd(dst, s8);
			} else if (lane == 1) {
				atomicAdd(dL_dopacity + id, s1);
			}
		}
	}
}

} // namespace moss
