// K1: per-Gaussian forward preprocess (+ per-tile instance counting), K8: mark-visible.
//
// Semantics follow DGR/cuda_rasterizer/forward.cu:155-256 (preprocessCUDA), :74-113 (computeCov2D),
// :118-152 (computeCov3D), :20-71 (computeColorFromSH) and DGR/cuda_rasterizer/auxiliary.h:41-56,139-164
// (ndc2Pix in fp64, getRect, in_frustum).  The arithmetic is written with explicitly-rounded intrinsics in
// the order nvcc contracts the reference's GLM expressions, because radii / rects / depth bits must be
// bit-identical (they define the sorted key lists).  What is new here: one fused pass also bins the
// Gaussian into its tiles (atomic per-tile counters) so no per-Gaussian prefix scan, key duplication pass
// or global radix sort is needed downstream, and colour+depth are packed into one float4 per Gaussian.
#pragma once
#include "common.cuh"
#include "skin.cuh"

namespace moss {

// SH basis constants (real spherical harmonics up to degree 3), same values as
// DGR/cuda_rasterizer/auxiliary.h:22-39 / utils/sh_utils.py:24-55.
#define MOSS_SH_C0 0.28209479177387814f
#define MOSS_SH_C1 0.4886025119029199f
#define MOSS_SH_C2_0 1.0925484305920792f
#define MOSS_SH_C2_1 -1.0925484305920792f
#define MOSS_SH_C2_2 0.31539156525252005f
#define MOSS_SH_C2_3 -1.0925484305920792f
#define MOSS_SH_C2_4 0.5462742152960396f
#define MOSS_SH_C3_0 -0.5900435899266435f
#define MOSS_SH_C3_1 2.890611442640554f
#define MOSS_SH_C3_2 -0.4570457994644658f
#define MOSS_SH_C3_3 0.3731763325901154f
#define MOSS_SH_C3_4 -0.4570457994644658f
#define MOSS_SH_C3_5 1.445305721320277f
#define MOSS_SH_C3_6 -0.5900435899266435f

struct Float3 { float x, y, z; };

// 3D covariance from scale and UN-normalised quaternion (r,x,y,z) -- DGR forward.cu:118-152.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod,
	float r, float x, float y, float z, float* cov)
{
	const float s0 = fmul(mod, sx), s1 = fmul(mod, sy), s2 = fmul(mod, sz);
	// Rounding sequence read off the reference's sm_100 SASS (ptxas fuses one product of every
	// "a*b +- c*d" pair that NVVM left unfused): e.g. xy - rz = fma(x, y, -(r*z)).
	const float yy = fmul(y, y), zz = fmul(z, z);
	const float xz = fmul(x, z), rx = fmul(r, x), rz = fmul(r, z);
	const float yy_zz = fadd(yy, zz);
	const float xx_zz = ffma(x, x, zz);
	const float xx_yy = ffma(x, x, yy);
	const float xy_m_rz = ffma(x, y, -rz), xy_p_rz = ffma(x, y, rz);
	const float ry_p_xz = ffma(r, y, xz), xz_m_ry = ffma(-r, y, xz);
	const float yz_m_rx = ffma(y, z, -rx), rx_p_yz = ffma(y, z, rx);
	// R listed the way the reference hands it to glm::mat3 (column i, row j)
	const float R00 = fsub(1.f, fadd(yy_zz, yy_zz));
	const float R01 = fadd(xy_m_rz, xy_m_rz);
	const float R02 = fadd(ry_p_xz, ry_p_xz);
	const float R10 = fadd(xy_p_rz, xy_p_rz);
	const float R11 = fsub(1.f, fadd(xx_zz, xx_zz));
	const float R12 = fadd(yz_m_rx, yz_m_rx);
	const float R20 = fadd(xz_m_ry, xz_m_ry);
	const float R21 = fadd(rx_p_yz, rx_p_yz);
	const float R22 = fsub(1.f, fadd(xx_yy, xx_yy));
	// M = S * R (S diagonal): M[i][j] = s_j * R[i][j]
	const float M00 = fmul(s0, R00), M01 = fmul(s1, R01), M02 = fmul(s2, R02);
	const float M10 = fmul(s0, R10), M11 = fmul(s1, R11), M12 = fmul(s2, R12);
	const float M20 = fmul(s0, R20), M21 = fmul(s1, R21), M22 = fmul(s2, R22);
	// Sigma = M^T M
	cov[0] = dot3_glm(M00, M00, M01, M01, M02, M02);
	cov[1] = dot3_glm(M10, M00, M11, M01, M12, M02);
	cov[2] = dot3_glm(M20, M00, M21, M01, M22, M02);
	cov[3] = dot3_glm(M10, M10, M11, M11, M12, M12);
	cov[4] = dot3_glm(M20, M10, M21, M11, M22, M12);
	cov[5] = dot3_glm(M20, M20, M21, M21, M22, M22);
}

// EWA projection of the 3D covariance: returns (a, b, c) of the 2D covariance incl. the 0.3 px low-pass.
// DGR forward.cu:74-113.  `v` = view matrix staged in shared memory.
__device__ __forceinline__ Float3 cov2d_ewa(float tx, float ty, float tz, const float* v,
	float focal_x, float focal_y, float tanfovx, float tanfovy, const float* c3, float* Tout = nullptr)
{
	const float limx = fmul(1.3f, tanfovx), limy = fmul(1.3f, tanfovy);
	const float txtz = fdiv(tx, tz), tytz = fdiv(ty, tz);
	const float cx = fminf(limx, fmaxf(-limx, txtz));
	const float cy = fminf(limy, fmaxf(-limy, tytz));
	const float tz2 = fmul(tz, tz);
	const float J00 = fdiv(focal_x, tz);
	const float J02 = fdiv(-fmul(focal_x, fmul(cx, tz)), tz2);
	const float J11 = fdiv(focal_y, tz);
	const float J12 = fdiv(-fmul(focal_y, fmul(cy, tz)), tz2);
	// T = W * J (GLM), only the two non-zero columns
	const float T00 = ffma(v[2], J02, fmul(v[0], J00));
	const float T01 = ffma(v[6], J02, fmul(v[4], J00));
	const float T02 = ffma(v[10], J02, fmul(v[8], J00));
	const float T10 = ffma(v[2], J12, fmul(J11, v[1]));
	const float T11 = ffma(v[6], J12, fmul(J11, v[5]));
	const float T12 = ffma(v[10], J12, fmul(J11, v[9]));
	if (Tout) { Tout[0] = T00; Tout[1] = T01; Tout[2] = T02; Tout[3] = T10; Tout[4] = T11; Tout[5] = T12; }
	// A = T^T * Vrk^T
	const float A00 = dot3_glm(T00, c3[0], T01, c3[1], T02, c3[2]);
	const float A01 = dot3_glm(T10, c3[0], T11, c3[1], T12, c3[2]);
	const float A10 = dot3_glm(T00, c3[1], T01, c3[3], T02, c3[4]);
	const float A11 = dot3_glm(T10, c3[1], T11, c3[3], T12, c3[4]);
	const float A20 = dot3_glm(T00, c3[2], T01, c3[4], T02, c3[5]);
	const float A21 = dot3_glm(T10, c3[2], T11, c3[4], T12, c3[5]);
	// cov = A * T
	Float3 r;
	r.x = fadd(dot3_glm(T00, A00, T01, A10, T02, A20), 0.3f);
	r.y = dot3_glm(T00, A01, T01, A11, T02, A21);
	r.z = fadd(dot3_glm(T10, A01, T11, A11, T12, A21), 0.3f);
	return r;
}

// SH (degree D) -> rgb, DGR forward.cu:20-71.  sh holds this Gaussian's coefficients (registers: every index below is a
// compile-time constant), loaded up front by sh_load.
__device__ __forceinline__ void sh_to_rgb(int D, const float (&sh)[48], float dx, float dy, float dz,
	float* rgb)
{
	const float len = fsqrt(ffma(dz, dz, ffma(dx, dx, fmul(dy, dy))));
	const float x = fdiv(dx, len), y = fdiv(dy, len), z = fdiv(dz, len);
#pragma unroll
	for (int c = 0; c < 3; c++)
		rgb[c] = fmul(sh[c], MOSS_SH_C0);
	if (D > 0) {
		const float a1 = fmul(y, MOSS_SH_C1), a2 = fmul(z, MOSS_SH_C1), a3 = fmul(x, MOSS_SH_C1);
#pragma unroll
		for (int c = 0; c < 3; c++) {
			float r = ffma(-a1, sh[3 + c], rgb[c]);
			r = ffma(a2, sh[6 + c], r);
			rgb[c] = ffma(-a3, sh[9 + c], r);
		}
		if (D > 1) {
			const float xx = fmul(x, x), yy = fmul(y, y), zz = fmul(z, z);
			const float xy = fmul(x, y), yz = fmul(y, z), xz = fmul(x, z);
			const float zz2 = fadd(zz, zz);
			const float xx_yy = fsub(xx, yy);
			const float b4 = fmul(xy, MOSS_SH_C2_0);
			const float b5 = fmul(yz, MOSS_SH_C2_1);
			const float b6 = fmul(fsub(fsub(zz2, xx), yy), MOSS_SH_C2_2);
			const float b7 = fmul(xz, MOSS_SH_C2_3);
			const float b8 = fmul(xx_yy, MOSS_SH_C2_4);
#pragma unroll
			for (int c = 0; c < 3; c++) {
				float r = ffma(b4, sh[12 + c], rgb[c]);
				r = ffma(b5, sh[15 + c], r);
				r = ffma(b6, sh[18 + c], r);
				r = ffma(b7, sh[21 + c], r);
				rgb[c] = ffma(b8, sh[24 + c], r);
			}
			if (D > 2) {
				// fused forms as in the reference's SASS: 3xx-yy = fma(xx,3,-yy); 4zz-xx-yy = fma(zz,4,-xx)-yy;
				// 2zz-3xx-3yy = fma(yy,-3, fma(xx,-3, 2zz)); xx-3yy = fma(yy,-3,xx)
				const float zz4_xx_yy = fsub(ffma(zz, 4.f, -xx), yy);
				const float c9 = fmul(fmul(y, MOSS_SH_C3_0), ffma(xx, 3.f, -yy));
				const float c10 = fmul(fmul(xy, MOSS_SH_C3_1), z);
				const float c11 = fmul(fmul(y, MOSS_SH_C3_2), zz4_xx_yy);
				const float c12 = fmul(fmul(z, MOSS_SH_C3_3), ffma(yy, -3.f, ffma(xx, -3.f, zz2)));
				const float c13 = fmul(fmul(x, MOSS_SH_C3_4), zz4_xx_yy);
				const float c14 = fmul(fmul(z, MOSS_SH_C3_5), xx_yy);
				const float c15 = fmul(fmul(x, MOSS_SH_C3_6), ffma(yy, -3.f, xx));
#pragma unroll
				for (int c = 0; c < 3; c++) {
					float r = ffma(c9, sh[27 + c], rgb[c]);
					r = ffma(c10, sh[30 + c], r);
					r = ffma(c11, sh[33 + c], r);
					r = ffma(c12, sh[36 + c], r);
					r = ffma(c13, sh[39 + c], r);
					r = ffma(c14, sh[42 + c], r);
					rgb[c] = ffma(c15, sh[45 + c], r);
				}
			}
		}
	}
#pragma unroll
	for (int c = 0; c < 3; c++)
		rgb[c] = fadd(rgb[c], 0.5f);
}

// The 3(D+1)^2 coefficients of one Gaussian as 16 B vector loads (the [M,3] block is 192 B for M = 16): 12 load
// instructions per thread instead of 48 scalar ones -- every lane touches its own sectors, so the LSU cost is per
// instruction.  Unaligned views fall back to scalar loads.
__device__ __forceinline__ void sh_load(int D, int M, const float* __restrict__ row, bool vec_ok, float (&sh)[48])
{
	const int nfl = 3 * (D + 1) * (D + 1);
	if (vec_ok) {
		const float4* r4 = reinterpret_cast<const float4*>(row);
		const int nvec = (nfl + 3) >> 2;      // never past the row: 3M floats with 3M % 4 == 0 and nfl <= 3M
#pragma unroll
		for (int k = 0; k < 12; k++) {
			float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
			if (k < nvec) v = __ldg(r4 + k);
			sh[4 * k] = v.x; sh[4 * k + 1] = v.y; sh[4 * k + 2] = v.z; sh[4 * k + 3] = v.w;
		}
	} else {
#pragma unroll
		for (int k = 0; k < 48; k++) sh[k] = k < nfl ? __ldg(row + k) : 0.f;
	}
}

// One quaternion row: a 16 B load when the array is 16 B-aligned (torch allocations are), four scalar loads for a
// view that is not (e.g. a slice of a packed buffer whose offset is not a multiple of 4 floats).
__device__ __forceinline__ float4 load_quat(const float* __restrict__ rotations, int i)
{
	if ((reinterpret_cast<uintptr_t>(rotations) & 15) == 0) return reinterpret_cast<const float4*>(rotations)[i];
	return make_float4(rotations[4 * i], rotations[4 * i + 1], rotations[4 * i + 2], rotations[4 * i + 3]);
}
__device__ __forceinline__ void store_quat(float* __restrict__ out, int i, float4 v)
{
	if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) { reinterpret_cast<float4*>(out)[i] = v; return; }
	out[4 * i] = v.x; out[4 * i + 1] = v.y; out[4 * i + 2] = v.z; out[4 * i + 3] = v.w;
}

// ndc -> pixel, evaluated in fp64 like the reference (DGR auxiliary.h:41-44).
__device__ __forceinline__ float ndc_to_pix(float v, int S)
{
	return (float)(__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5));
}

constexpr int PRE_THREADS = 128;   // 20 k Gaussians -> 157 blocks: every SM gets work (256 left 69 of 148 idle)

// What the fused skin + preprocess entry points add to the per-Gaussian kernels (SKIN = true; north_star subsystem 1,
// SURVEY.md section 7 step 8): the Gaussian's world position and covariance come out of the skinning device functions
// (skin.cuh) in registers instead of being read from HBM -- MOSS's one-view-per-iteration loop
// (gaussian_renderer/__init__.py:57-93 feeding DGR forward.cu:155-256) never needs them in memory.
struct SkinFused {
	SkinArgs a;
	float* means3D;      // forward: optional [P,3] copy of the skinned centres (NULL: never written)
	float* transforms;   // forward: optional [P,9] (render()'s "transforms")
};

// grid = (ceil(P / PRE_THREADS), V): block row v preprocesses the Gaussians for view v ("virtual Gaussian" v*P + i).
#ifndef MOSS_PRE_FWD_MAXREG
#define MOSS_PRE_FWD_MAXREG 0
#endif
#if MOSS_PRE_FWD_MAXREG > 0
#define MOSS_PRE_FWD_REGCAP __maxnreg__(SKIN ? 255 : MOSS_PRE_FWD_MAXREG)
#else
#define MOSS_PRE_FWD_REGCAP __launch_bounds__(PRE_THREADS)
#endif
template <bool SKIN>
__global__ void MOSS_PRE_FWD_REGCAP
preprocess_forward_kernel(const __grid_constant__ ViewBatch vb,
	const float* __restrict__ means3D, const float* __restrict__ shs, const float* __restrict__ colors_precomp,
	const float* __restrict__ opacities, const float* __restrict__ scales, const float* __restrict__ rotations,
	const float* __restrict__ cov3D_precomp, int prefiltered,
	GeomWS g, uint32_t* __restrict__ tile_count, int* __restrict__ radii, const SkinFused sk)
{
	// Every global read of the Gaussian is issued here, before the first data-dependent branch, so the kernel pays
	// ONE DRAM round trip instead of a chain of four (position -> covariance -> SH -> opacity); with a cold L2 and
	// one block per SM that chain was most of the launch's duration.  Culled Gaussians read a little more than needed.
	const int P = vb.P;
	const int view = blockIdx.y;
	const ViewCam& cam = vb.cam[view];
	const int idx = blockIdx.x * PRE_THREADS + threadIdx.x;
	const bool active = idx < P;
	const int li = active ? idx : 0;
	const size_t vid = (size_t)view * P + idx;            // virtual Gaussian: row of the per-view state
	float px, py, pz;
	const float opac = opacities[li];
	float c3[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
	float4 quat = make_float4(0.f, 0.f, 0.f, 0.f);
	float sc3[3] = {0.f, 0.f, 0.f};
	int sk_vid = 0;
	SkinOwn sk_own;
	if constexpr (SKIN) {
		sk_vid = __ldg(sk.a.vert_ids + li);
		sk_own = skin_own(sk.a, li);
		px = py = pz = 0.f;
	} else {
		px = means3D[3 * li]; py = means3D[3 * li + 1]; pz = means3D[3 * li + 2];
		if (cov3D_precomp != nullptr) {
#pragma unroll
			for (int k = 0; k < 6; k++) c3[k] = cov3D_precomp[6 * li + k];
		} else {
			quat = load_quat(rotations, li);
			sc3[0] = scales[3 * li]; sc3[1] = scales[3 * li + 1]; sc3[2] = scales[3 * li + 2];
		}
	}
	float shl[48];
	float rgb_in[3] = {0.f, 0.f, 0.f};
	if (colors_precomp == nullptr) {
		const bool vec_ok = ((reinterpret_cast<uintptr_t>(shs) & 15) == 0) && ((3 * vb.M) % 4 == 0);
		sh_load(vb.D, vb.M, shs + (size_t)li * vb.M * 3, vec_ok, shl);
	} else {
#pragma unroll
		for (int c = 0; c < 3; c++) rgb_in[c] = colors_precomp[3 * li + c];
	}

	__shared__ float s_view[16], s_proj[16], s_cam[3];
	if (threadIdx.x < 16) s_view[threadIdx.x] = cam.view[threadIdx.x];
	else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = cam.proj[threadIdx.x - 16];
	else if (threadIdx.x < 35 && colors_precomp == nullptr) s_cam[threadIdx.x - 32] = cam.campos[threadIdx.x - 32];
	__syncthreads();
	if constexpr (SKIN) {
		// skinning in registers: blended joint transforms, offsets, world position, transformed covariance
		__shared__ SkinShared sh;
		skin_stage(sh, sk.a);
		if (active) {
			float bw[NJ];
			skin_weights(sk.a, li, sk_vid, bw);
			SkinFwd f;
			skin_point(sh, sk.a, sk_own.x, sk_vid, bw, f);
			float world[3];
			skin_world(sh, f, world);
			px = world[0]; py = world[1]; pz = world[2];
			skin_cov(sk_own, f, c3);
			if (sk.means3D) { sk.means3D[3 * (size_t)li] = px; sk.means3D[3 * (size_t)li + 1] = py; sk.means3D[3 * (size_t)li + 2] = pz; }
			if (sk.transforms) {
#pragma unroll
				for (int i = 0; i < 9; i++) sk.transforms[9 * (size_t)li + i] = f.Tr.m[i];
			}
		}
	}
	if (!active) return;

	int my_radius = 0;
	uint32_t touched = 0;
	const float vz = affine_row(px, py, pz, s_view[2], s_view[6], s_view[10], s_view[14]);

	do {
		if (vz <= 0.2f) {  // near cull, DGR auxiliary.h:154
			if (prefiltered) {
				printf("Point is filtered although prefiltered is set. This shouldn't happen!");
				__trap();
			}
			break;
		}
		const float hx = affine_row(px, py, pz, s_proj[0], s_proj[4], s_proj[8], s_proj[12]);
		const float hy = affine_row(px, py, pz, s_proj[1], s_proj[5], s_proj[9], s_proj[13]);
		const float hw = affine_row(px, py, pz, s_proj[3], s_proj[7], s_proj[11], s_proj[15]);
		const float p_w = frcp(fadd(hw, 0.0000001f));
		const float ndc_x = fmul(hx, p_w), ndc_y = fmul(hy, p_w);

		if constexpr (!SKIN) {
			if (cov3D_precomp == nullptr) {
				cov3d_from_scale_rot(sc3[0], sc3[1], sc3[2], vb.scale_modifier, quat.x, quat.y, quat.z, quat.w, c3);
#pragma unroll
				for (int k = 0; k < 6; k++) g.cov3D[6 * vid + k] = c3[k];
			}
		}

		const float tx = affine_row(px, py, pz, s_view[0], s_view[4], s_view[8], s_view[12]);
		const float ty = affine_row(px, py, pz, s_view[1], s_view[5], s_view[9], s_view[13]);
		const Float3 cov = cov2d_ewa(tx, ty, vz, s_view, cam.focal_x, cam.focal_y, cam.tanfovx, cam.tanfovy, c3);

		const float det = ffma(cov.x, cov.z, -fmul(cov.y, cov.y));   // SASS: FMUL b*b; FFMA a*c-bb
		if (det == 0.0f) break;
		const float det_inv = frcp(det);
		const float conic_x = fmul(cov.z, det_inv);
		const float conic_y = fmul(det_inv, -cov.y);
		const float conic_z = fmul(cov.x, det_inv);

		const float mid = fmul(fadd(cov.x, cov.z), 0.5f);
		const float sq = fsqrt(fmaxf(ffma(mid, mid, -det), 0.1f));
		const float lmax = fmaxf(fadd(mid, sq), fsub(mid, sq));
		const float radius_f = ceilf(fmul(fsqrt(lmax), 3.f));
		const float pix_x = ndc_to_pix(ndc_x, vb.W), pix_y = ndc_to_pix(ndc_y, vb.H);
		const int ri = (int)radius_f;              // DGR auxiliary.h:46 (float -> int max_radius)
		// The low-pass filter keeps every finite covariance at radius >= 2; 0 means a NaN covariance (float -> int of NaN).
		// The reference still counts such a Gaussian in num_rendered but never emits its key (radii > 0 test in
		// duplicateWithKeys, DGR rasterizer_impl.cu:82), so its sorted list holds an UNINITIALISED entry that the blend
		// kernels then use as a Gaussian index.  Here it is culled like any other invisible Gaussian.
		if (ri <= 0) break;
		const float rf = (float)ri;
		const uint32_t gx = vb.grid_x, gy = vb.grid_y;
		const uint32_t rminx = min(gx, (uint32_t)max(0, (int)fmul(fsub(pix_x, rf), 0.0625f)));
		const uint32_t rminy = min(gy, (uint32_t)max(0, (int)fmul(fsub(pix_y, rf), 0.0625f)));
		const uint32_t rmaxx = min(gx, (uint32_t)max(0, (int)fmul(fadd(fadd(fadd(pix_x, rf), 16.f), -1.f), 0.0625f)));
		const uint32_t rmaxy = min(gy, (uint32_t)max(0, (int)fmul(fadd(fadd(fadd(pix_y, rf), 16.f), -1.f), 0.0625f)));
		const uint32_t area = (rmaxx - rminx) * (rmaxy - rminy);
		if (area == 0) break;

		float rgb[3];
		uint8_t clamp_bits = 0;
		if (colors_precomp == nullptr) {
			sh_to_rgb(vb.D, shl, fsub(px, s_cam[0]), fsub(py, s_cam[1]), fsub(pz, s_cam[2]), rgb);
#pragma unroll
			for (int c = 0; c < 3; c++) {
				if (rgb[c] < 0.f) { clamp_bits |= (1u << c); rgb[c] = 0.f; }
			}
		} else {
#pragma unroll
			for (int c = 0; c < 3; c++) rgb[c] = rgb_in[c];
		}

		g.depths[vid] = vz;
		g.means2D[vid] = make_float2(pix_x, pix_y);
		g.conic_opacity[vid] = make_float4(conic_x, conic_y, conic_z, opac);
		g.feat[vid] = make_float4(rgb[0], rgb[1], rgb[2], vz);
		g.clamped[vid] = clamp_bits;
		g.rect[vid] = make_ushort4((unsigned short)rminx, (unsigned short)rminy, (unsigned short)rmaxx, (unsigned short)rmaxy);
		// the blend backward accumulates into this row; culled Gaussians never get one (and are never read back)
#pragma unroll
		for (int k = 0; k < GRAD_ACC_VEC; k++) g.grad_acc[GRAD_ACC_VEC * vid + k] = make_float4(0.f, 0.f, 0.f, 0.f);
		my_radius = ri;
		touched = area;
		// bin: one counter per tile (replaces the per-Gaussian scan + key duplication + global sort)
		const uint32_t sub = ((uint32_t)idx >> 5) & (TILE_SUB - 1);      // one sub-counter per counting warp
		uint32_t* tc = tile_count + (size_t)view * vb.T * TILE_SUB;
		for (uint32_t y = rminy; y < rmaxy; y++)
			for (uint32_t x = rminx; x < rmaxx; x++)
				atomicAdd(&tc[(y * gx + x) * TILE_SUB + sub], 1u);
	} while (false);

	radii[vid] = my_radius;
	g.tiles_touched[vid] = touched;
}

// present[i] = view.z > 0.2   (DGR rasterizer_impl.cu:54-66 with auxiliary.h:139-164)
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
	uint8_t* __restrict__ present)
{
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= P) return;
	const float vz = affine_row(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2],
		view[2], view[6], view[10], view[14]);
	present[idx] = !(vz <= 0.2f);
}

} // namespace moss
