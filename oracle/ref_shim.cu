// TEST INFRASTRUCTURE ONLY -- never linked into or loaded by the product (moss_b200).
//
// C-ABI shim (our code) around the UNMODIFIED reference rasterizer
//   DGR/cuda_rasterizer/rasterizer.h:24-89  (CudaRasterizer::Rasterizer::forward/backward/markVisible)
// compiled together with the reference's forward.cu / backward.cu / rasterizer_impl.cu
// by oracle/build_ref.py into oracle/_ref/libref_raster.so.  It owns the three growable
// state chunks the reference asks for through std::function callbacks
// (DGR/rasterize_points.cu:27-33 does the same with torch tensors) and re-derives the
// state pointers with the reference's own fromChunk (DGR/cuda_rasterizer/rasterizer_impl.cu:155-194)
// so that parity tests can compare internal state bit for bit.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <functional>
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"

namespace {
struct Chunk {
	char* ptr = nullptr;
	size_t cap = 0;
	char* grow(size_t n) {
		if (n > cap) {
			if (ptr) cudaFree(ptr);
			cap = n + (n >> 2) + 1024;
			cudaMalloc(&ptr, cap);
		}
		return ptr;
	}
};
Chunk g_geom, g_bin, g_img;
int g_P = 0, g_R = 0, g_W = 0, g_H = 0;
}

extern "C" {

int ref_forward(int P, int D, int M, const float* bg, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tanfovx, float tanfovy, int prefiltered,
	float* out_color, float* out_depth, float* out_alpha, int* radii, int debug)
{
	std::function<char*(size_t)> fg = [](size_t n) { return g_geom.grow(n); };
	std::function<char*(size_t)> fb = [](size_t n) { return g_bin.grow(n); };
	std::function<char*(size_t)> fi = [](size_t n) { return g_img.grow(n); };
	g_P = P; g_W = W; g_H = H;
	g_R = CudaRasterizer::Rasterizer::forward(fg, fb, fi, P, D, M, bg, W, H, means3D, shs, colors_precomp,
		opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
		tanfovx, tanfovy, prefiltered != 0, out_color, out_depth, out_alpha, radii, debug != 0);
	return g_R;
}

void ref_backward(int P, int D, int M, int R, const float* bg, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* alphas,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tanfovx, float tanfovy, const int* radii,
	const float* dL_dpix, const float* dL_ddepths, const float* dL_dalphas,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
	float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int debug)
{
	CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, alphas,
		scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
		radii, g_geom.ptr, g_bin.ptr, g_img.ptr, dL_dpix, dL_ddepths, dL_dalphas,
		dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
		debug != 0);
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present)
{
	CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

// which: 0 depths f32[P] | 1 clamped u8[3P] | 2 means2D f32[2P] | 3 cov3D f32[6P] | 4 conic_opacity f32[4P]
//        5 rgb f32[3P] | 6 point_offsets u32[P] | 7 tiles_touched u32[P]
//        8 keys_unsorted u64[R] | 9 keys u64[R] | 10 list_unsorted u32[R] | 11 list u32[R]
//        12 ranges u32[2*tiles] | 13 n_contrib u32[N] | 14 accum_alpha f32[N]
// Copies the array into dst (device pointer); returns its size in bytes (or -1).
long long ref_export(int which, void* dst, long long max_bytes)
{
	using namespace CudaRasterizer;
	char* cg_ = g_geom.ptr; char* cb_ = g_bin.ptr; char* ci_ = g_img.ptr;
	const void* src = nullptr; long long n = -1;
	size_t P = g_P, R = g_R, N = (size_t)g_W * g_H;
	size_t tiles = (size_t)((g_W + 15) / 16) * ((g_H + 15) / 16);
	if (which <= 7) {
		if (!cg_) return -1;
		GeometryState s = GeometryState::fromChunk(cg_, P);
		switch (which) {
		case 0: src = s.depths; n = 4 * P; break;
		case 1: src = s.clamped; n = 3 * P; break;
		case 2: src = s.means2D; n = 8 * P; break;
		case 3: src = s.cov3D; n = 24 * P; break;
		case 4: src = s.conic_opacity; n = 16 * P; break;
		case 5: src = s.rgb; n = 12 * P; break;
		case 6: src = s.point_offsets; n = 4 * P; break;
		case 7: src = s.tiles_touched; n = 4 * P; break;
		}
	} else if (which <= 11) {
		if (!cb_) return -1;
		BinningState s = BinningState::fromChunk(cb_, R);
		switch (which) {
		case 8: src = s.point_list_keys_unsorted; n = 8 * R; break;
		case 9: src = s.point_list_keys; n = 8 * R; break;
		case 10: src = s.point_list_unsorted; n = 4 * R; break;
		case 11: src = s.point_list; n = 4 * R; break;
		}
	} else {
		if (!ci_) return -1;
		ImageState s = ImageState::fromChunk(ci_, N);
		switch (which) {
		case 12: src = s.ranges; n = 8 * tiles; break;
		case 13: src = s.n_contrib; n = 4 * N; break;
		case 14: src = s.accum_alpha; n = 4 * N; break;
		}
	}
	if (n < 0 || n > max_bytes) return -1;
	if (n > 0) cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice);
	return n;
}

int ref_last_cuda_error() { return (int)cudaDeviceSynchronize(); }

}
