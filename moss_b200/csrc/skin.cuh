// Fused per-Gaussian skinning + covariance front end.
//
// One thread per Gaussian does what MOSS's render() does between pc.get_xyz and the rasterizer call
// (gaussian_renderer/__init__.py:57-85) in ~60 small torch kernels:
//   GaussianModel.coarse_deform_c2source, per-Gaussian half   scene/gaussian_model.py:829-846, 857-868, 889-921
//   build_covariance_from_scaling_rotation                    scene/gaussian_model.py:37-44
//   build_rotation / build_scaling_rotation / strip_symmetric utils/general_utils.py:62-118
// The two 24-joint transform tables (big pose, target pose) are staged once per block in shared memory; the
// per-Gaussian reads are one 96 B weight row (6 x 16 B), three 12 B offset rows and the Gaussian's own 40 B.
#pragma once
#include "common.cuh"

namespace moss {

constexpr int SKIN_THREADS = 128;
constexpr int NJ = 24;            // SMPL joints (smplx/lbs.py, nets/mlp_delta_body_pose.py:41)

struct SkinArgs {
	int P;
	const float* xyz;          // [P,3] canonical (big-pose) centres
	const int*   vert_ids;     // [P]   nearest SMPL vertex (scene/gaussian_model.py:827)
	const float* weights;      // [V,24] SMPL skinning weights
	const float* lbs_delta;    // [P,24] or NULL: bw = softmax(log(w + 1e-9) + delta)  (:831-833)
	const float* A_big;        // [24,16] row-major 4x4, big pose -> T pose   (:838)
	const float* A_tgt;        // [24,16] T pose -> target pose               (:895)
	const float* poseoff_big;  // [V,3] or NULL  (:849-857)
	const float* shapeoff;     // [V,3] or NULL  (:862-866)
	const float* poseoff_tgt;  // [V,3] or NULL  (:872-889)
	const float* Rg;           // [9]  global rotation params['R']
	const float* Th;           // [3]
	const float* scaling;      // [P,3] activated scales, or NULL (no covariance)
	const float* rotation;     // [P,4] raw quaternion (r,x,y,z); normalised here (utils/general_utils.py:79-81)
	float scale_modifier;
};

struct Mat3 { float m[9]; };   // row-major

__device__ __forceinline__ Mat3 mat3_mul(const Mat3& a, const Mat3& b)
{
	Mat3 r;
#pragma unroll
	for (int i = 0; i < 3; i++)
#pragma unroll
		for (int j = 0; j < 3; j++)
			r.m[3 * i + j] = a.m[3 * i] * b.m[j] + a.m[3 * i + 1] * b.m[3 + j] + a.m[3 * i + 2] * b.m[6 + j];
	return r;
}
__device__ __forceinline__ Mat3 mat3_mul_tn(const Mat3& a, const Mat3& b)   // a^T b
{
	Mat3 r;
#pragma unroll
	for (int i = 0; i < 3; i++)
#pragma unroll
		for (int j = 0; j < 3; j++)
			r.m[3 * i + j] = a.m[i] * b.m[j] + a.m[3 + i] * b.m[3 + j] + a.m[6 + i] * b.m[6 + j];
	return r;
}
__device__ __forceinline__ Mat3 mat3_mul_nt(const Mat3& a, const Mat3& b)   // a b^T
{
	Mat3 r;
#pragma unroll
	for (int i = 0; i < 3; i++)
#pragma unroll
		for (int j = 0; j < 3; j++)
			r.m[3 * i + j] = a.m[3 * i] * b.m[3 * j] + a.m[3 * i + 1] * b.m[3 * j + 1] + a.m[3 * i + 2] * b.m[3 * j + 2];
	return r;
}
__device__ __forceinline__ Mat3 mat3_inverse(const Mat3& a)
{
	const float c00 = a.m[4] * a.m[8] - a.m[5] * a.m[7];
	const float c01 = a.m[5] * a.m[6] - a.m[3] * a.m[8];
	const float c02 = a.m[3] * a.m[7] - a.m[4] * a.m[6];
	const float det = a.m[0] * c00 + a.m[1] * c01 + a.m[2] * c02;
	const float id = 1.0f / det;
	Mat3 r;
	r.m[0] = c00 * id; r.m[1] = (a.m[2] * a.m[7] - a.m[1] * a.m[8]) * id; r.m[2] = (a.m[1] * a.m[5] - a.m[2] * a.m[4]) * id;
	r.m[3] = c01 * id; r.m[4] = (a.m[0] * a.m[8] - a.m[2] * a.m[6]) * id; r.m[5] = (a.m[2] * a.m[3] - a.m[0] * a.m[5]) * id;
	r.m[6] = c02 * id; r.m[7] = (a.m[1] * a.m[6] - a.m[0] * a.m[7]) * id; r.m[8] = (a.m[0] * a.m[4] - a.m[1] * a.m[3]) * id;
	return r;
}
__device__ __forceinline__ void mat3_vec(const Mat3& a, const float v[3], float out[3])
{
#pragma unroll
	for (int i = 0; i < 3; i++) out[i] = a.m[3 * i] * v[0] + a.m[3 * i + 1] * v[1] + a.m[3 * i + 2] * v[2];
}
__device__ __forceinline__ void mat3_tvec(const Mat3& a, const float v[3], float out[3])  // a^T v
{
#pragma unroll
	for (int i = 0; i < 3; i++) out[i] = a.m[i] * v[0] + a.m[3 + i] * v[1] + a.m[6 + i] * v[2];
}

// Stage the top three rows of both joint tables (2 x 24 x 12 floats) and the global transform in shared memory.
struct SkinShared {
	float Ab[NJ][12];
	float At[NJ][12];
	float Rg[9], Rgi[9], Th[3];
};
__device__ __forceinline__ void skin_stage(SkinShared& sh, const SkinArgs& a)
{
	for (int i = threadIdx.x; i < NJ * 12; i += blockDim.x) {
		const int j = i / 12, e = i % 12;
		sh.Ab[j][e] = a.A_big[j * 16 + e];
		sh.At[j][e] = a.A_tgt[j * 16 + e];
	}
	if (threadIdx.x == 0) {
		Mat3 R;
		for (int i = 0; i < 9; i++) { R.m[i] = a.Rg[i]; sh.Rg[i] = R.m[i]; }
		const Mat3 Ri = mat3_inverse(R);          // torch.inverse(R), scene/gaussian_model.py:911
		for (int i = 0; i < 9; i++) sh.Rgi[i] = Ri.m[i];
		for (int i = 0; i < 3; i++) sh.Th[i] = a.Th[i];
	}
	__syncthreads();
}

// Blend weights of one Gaussian: table row, optionally re-weighted by softmax(log(w + 1e-9) + delta).
__device__ __forceinline__ void skin_weights(const SkinArgs& a, int idx, int vid, float bw[NJ])
{
	const float4* row = reinterpret_cast<const float4*>(a.weights + (size_t)vid * NJ);
#pragma unroll
	for (int k = 0; k < NJ / 4; k++) {
		const float4 w = __ldg(row + k);
		bw[4 * k] = w.x; bw[4 * k + 1] = w.y; bw[4 * k + 2] = w.z; bw[4 * k + 3] = w.w;
	}
	if (a.lbs_delta) {
		const float4* drow = reinterpret_cast<const float4*>(a.lbs_delta + (size_t)idx * NJ);
		float mx = -INFINITY;
#pragma unroll
		for (int k = 0; k < NJ / 4; k++) {
			const float4 d = __ldg(drow + k);
			bw[4 * k] = logf(bw[4 * k] + 1e-9f) + d.x; bw[4 * k + 1] = logf(bw[4 * k + 1] + 1e-9f) + d.y;
			bw[4 * k + 2] = logf(bw[4 * k + 2] + 1e-9f) + d.z; bw[4 * k + 3] = logf(bw[4 * k + 3] + 1e-9f) + d.w;
		}
#pragma unroll
		for (int j = 0; j < NJ; j++) mx = fmaxf(mx, bw[j]);
		float sum = 0.f;
#pragma unroll
		for (int j = 0; j < NJ; j++) { bw[j] = expf(bw[j] - mx); sum += bw[j]; }
		const float inv = 1.0f / sum;
#pragma unroll
		for (int j = 0; j < NJ; j++) bw[j] *= inv;
	}
}

__device__ __forceinline__ void skin_blend(const float (*A)[12], const float bw[NJ], Mat3& R, float t[3])
{
	float acc[12];
#pragma unroll
	for (int e = 0; e < 12; e++) acc[e] = 0.f;
#pragma unroll
	for (int j = 0; j < NJ; j++) {
		const float w = bw[j];
		if (w != 0.f) {
#pragma unroll
			for (int e = 0; e < 12; e++) acc[e] = fmaf(w, A[j][e], acc[e]);
		}
	}
#pragma unroll
	for (int i = 0; i < 3; i++) {
		R.m[3 * i] = acc[4 * i]; R.m[3 * i + 1] = acc[4 * i + 1]; R.m[3 * i + 2] = acc[4 * i + 2];
		t[i] = acc[4 * i + 3];
	}
}

__device__ __forceinline__ void load3(const float* base, int row, float v[3])
{
	if (base) { v[0] = __ldg(base + 3 * (size_t)row); v[1] = __ldg(base + 3 * (size_t)row + 1); v[2] = __ldg(base + 3 * (size_t)row + 2); }
	else { v[0] = v[1] = v[2] = 0.f; }
}

// The Gaussian's own values, read at the top of the kernels together with vert_ids: the weight / offset rows depend on
// vert_ids (one extra round trip), nothing else should queue up behind them.
struct SkinOwn { float x[3], qr[4], sc[3]; };
__device__ __forceinline__ SkinOwn skin_own(const SkinArgs& a, int idx)
{
	SkinOwn o;
#pragma unroll
	for (int i = 0; i < 3; i++) o.x[i] = __ldg(a.xyz + 3 * (size_t)idx + i);
#pragma unroll
	for (int i = 0; i < 4; i++) o.qr[i] = a.rotation ? __ldg(a.rotation + 4 * (size_t)idx + i) : 0.f;
#pragma unroll
	for (int i = 0; i < 3; i++) o.sc[i] = a.scaling ? a.scale_modifier * __ldg(a.scaling + 3 * (size_t)idx + i) : 0.f;
	return o;
}

// Everything of the forward the backward needs again (recomputed there instead of being stored).
struct SkinFwd {
	Mat3 Rbi, Rt, Tr0, Tr;     // inverse big-pose rotation, target rotation, Rt*Rbi, Rg*Rt*Rbi
	float xb[3];               // x - t_big
	float q[3];                // T-pose point after the three offsets
	float s[3];                // SMPL-space posed point
};
__device__ __forceinline__ void skin_point(const SkinShared& sh, const SkinArgs& a, const float x[3], int vid, const float bw[NJ], SkinFwd& f)
{
	Mat3 Rb; float tb[3];
	skin_blend(sh.Ab, bw, Rb, tb);
	f.Rbi = mat3_inverse(Rb);
#pragma unroll
	for (int i = 0; i < 3; i++) f.xb[i] = x[i] - tb[i];
	float q0[3];
	mat3_vec(f.Rbi, f.xb, q0);
	float pb[3], so[3], pt[3];
	load3(a.poseoff_big, vid, pb); load3(a.shapeoff, vid, so); load3(a.poseoff_tgt, vid, pt);
#pragma unroll
	for (int i = 0; i < 3; i++) f.q[i] = ((q0[i] - pb[i]) + so[i]) + pt[i];
	float tt[3];
	skin_blend(sh.At, bw, f.Rt, tt);
	float rq[3];
	mat3_vec(f.Rt, f.q, rq);
#pragma unroll
	for (int i = 0; i < 3; i++) f.s[i] = rq[i] + tt[i];
	f.Tr0 = mat3_mul(f.Rt, f.Rbi);
	Mat3 Rg;
#pragma unroll
	for (int i = 0; i < 9; i++) Rg.m[i] = sh.Rg[i];
	f.Tr = mat3_mul(Rg, f.Tr0);
}

// q -> rotation matrix of the NORMALISED quaternion, L = R diag(mod * s) (utils/general_utils.py:78-118)
__device__ __forceinline__ void quat_rot(const float qr[4], float inv_norm, Mat3& R)
{
	const float r = qr[0] * inv_norm, x = qr[1] * inv_norm, y = qr[2] * inv_norm, z = qr[3] * inv_norm;
	R.m[0] = 1.f - 2.f * (y * y + z * z); R.m[1] = 2.f * (x * y - r * z); R.m[2] = 2.f * (x * z + r * y);
	R.m[3] = 2.f * (x * y + r * z); R.m[4] = 1.f - 2.f * (x * x + z * z); R.m[5] = 2.f * (y * z - r * x);
	R.m[6] = 2.f * (x * z - r * y); R.m[7] = 2.f * (y * z + r * x); R.m[8] = 1.f - 2.f * (x * x + y * y);
}

// world = smpl_pts @ inverse(R) + Th   (scene/gaussian_model.py:911-912)
__device__ __forceinline__ void skin_world(const SkinShared& sh, const SkinFwd& f, float world[3])
{
#pragma unroll
	for (int i = 0; i < 3; i++)
		world[i] = (f.s[0] * sh.Rgi[i] + f.s[1] * sh.Rgi[3 + i] + f.s[2] * sh.Rgi[6 + i]) + sh.Th[i];
}

// cov3D_precomp = strip_symmetric(T (L L^T) T^T), L = R(q / |q|) diag(mod * s)   (scene/gaussian_model.py:37-44)
__device__ __forceinline__ void skin_cov(const SkinOwn& own, const SkinFwd& f, float c6[6])
{
	const float (&qr)[4] = own.qr;
	const float (&sc)[3] = own.sc;
	const float inv_norm = 1.0f / sqrtf(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
	Mat3 Rq, L;
	quat_rot(qr, inv_norm, Rq);
#pragma unroll
	for (int i = 0; i < 3; i++)
#pragma unroll
		for (int k = 0; k < 3; k++) L.m[3 * i + k] = Rq.m[3 * i + k] * sc[k];
	const Mat3 S = mat3_mul_nt(L, L);              // L L^T
	const Mat3 TS = mat3_mul(f.Tr, S);             // transform @ cov
	const Mat3 C = mat3_mul_nt(TS, f.Tr);          // ... @ transform^T
	c6[0] = C.m[0]; c6[1] = C.m[1]; c6[2] = C.m[2]; c6[3] = C.m[4]; c6[4] = C.m[5]; c6[5] = C.m[8];   // upper triangle
}

__global__ void __launch_bounds__(SKIN_THREADS)
skin_forward_kernel(SkinArgs a, float* __restrict__ means3D, float* __restrict__ cov3D, float* __restrict__ transforms,
                    float* __restrict__ bweights, float* __restrict__ translation)
{
	__shared__ SkinShared sh;
	skin_stage(sh, a);
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= a.P) return;
	const int vid = __ldg(a.vert_ids + idx);
	const SkinOwn own = skin_own(a, idx);
	float bw[NJ];
	skin_weights(a, idx, vid, bw);
	SkinFwd f;
	skin_point(sh, a, own.x, vid, bw, f);
	float world[3];
	skin_world(sh, f, world);
#pragma unroll
	for (int i = 0; i < 3; i++) means3D[3 * (size_t)idx + i] = world[i];
	if (transforms) {
#pragma unroll
		for (int i = 0; i < 9; i++) transforms[9 * (size_t)idx + i] = f.Tr.m[i];
	}
	if (bweights) {
		float4* o = reinterpret_cast<float4*>(bweights + (size_t)idx * NJ);
#pragma unroll
		for (int k = 0; k < NJ / 4; k++) o[k] = make_float4(bw[4 * k], bw[4 * k + 1], bw[4 * k + 2], bw[4 * k + 3]);
	}
	if (translation) {
		// the same chain applied to the origin's image (return_transl, scene/gaussian_model.py:846-848, 858-868, 905-916)
		float tb[3]; Mat3 Rb;
		skin_blend(sh.Ab, bw, Rb, tb);
		float nt[3] = {-tb[0], -tb[1], -tb[2]}, tr[3];
		mat3_vec(f.Rbi, nt, tr);
		float pb[3], so[3], pt[3];
		load3(a.poseoff_big, vid, pb); load3(a.shapeoff, vid, so); load3(a.poseoff_tgt, vid, pt);
#pragma unroll
		for (int i = 0; i < 3; i++) tr[i] = ((tr[i] - pb[i]) + so[i]) + pt[i];
		Mat3 Rt; float tt[3], r2[3];
		skin_blend(sh.At, bw, Rt, tt);
		mat3_vec(Rt, tr, r2);
#pragma unroll
		for (int i = 0; i < 3; i++) r2[i] += tt[i];
#pragma unroll
		for (int i = 0; i < 3; i++)
			translation[3 * (size_t)idx + i] = (r2[0] * sh.Rgi[i] + r2[1] * sh.Rgi[3 + i] + r2[2] * sh.Rgi[6 + i]) + sh.Th[i];
	}
	if (cov3D && a.scaling) {
		float c6[6];
		skin_cov(own, f, c6);
		float* o = cov3D + 6 * (size_t)idx;
#pragma unroll
		for (int k = 0; k < 6; k++) o[k] = c6[k];
	}
}


// Backward of skin_forward_kernel (the reference's backward here is torch autograd through the python ops).
// Upstream: g_means3D [P,3] (from the rasterizer's dL/dmeans3D), g_cov3D [P,6] (dL/dcov3D_precomp; may be NULL).
// Written per Gaussian: d_xyz [P,3], d_scaling [P,3], d_rotation [P,4] (raw quaternion, includes the
// normalisation Jacobian), d_lbs_delta [P,24] (only with lbs_delta).  Accumulated (caller-zeroed):
// d_A_tgt [24,16] (row-major 4x4 per joint like A_tgt; bottom rows stay 0 -> pose-correction MLP) and
// d_poseoff_tgt [V,3] (scatter by vert_id).
// The per-Gaussian backward of the front end: upstream gw = dL/dmeans3D, gcv = dL/dcov3D (has_gcov: whether it is there).
// Writes d_xyz / d_scaling / d_rotation / d_lbs_delta rows, adds into d_poseoff_tgt (global atomics) and, with want_dA,
// hands back gAt_out[12] = the gradient w.r.t. this Gaussian's blended target transform (rows r00 r01 r02 t0 | ...): its
// contribution to dL/dA_tgt[j] is bw[j] * gAt_out, summed over the block by skin_accumulate_dA().  `f` and `bw` are the
// recomputed forward.
__device__ __forceinline__ void skin_backward_point(const SkinShared& sh, const SkinArgs& a, int idx, int vid, const SkinOwn& own,
	const float bw[NJ], const SkinFwd& f, const float gw[3], const float gcv[6], bool has_gcov,
	float* __restrict__ d_xyz, float* __restrict__ d_scaling, float* __restrict__ d_rotation, float* __restrict__ d_lbs_delta,
	float gAt_out[12], bool want_dA, float* __restrict__ d_poseoff_tgt)
{
	const bool d_A_tgt = want_dA;
	Mat3 dTr;
#pragma unroll
	for (int i = 0; i < 9; i++) dTr.m[i] = 0.f;
	if (a.scaling && has_gcov) {
		const float (&qr)[4] = own.qr;
		const float (&sc)[3] = own.sc;
		const float inv_norm = 1.0f / sqrtf(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
		Mat3 Rq, L;
		quat_rot(qr, inv_norm, Rq);
#pragma unroll
		for (int i = 0; i < 3; i++)
#pragma unroll
			for (int k = 0; k < 3; k++) L.m[3 * i + k] = Rq.m[3 * i + k] * sc[k];
		const Mat3 S = mat3_mul_nt(L, L);
		Mat3 G;      // gradient w.r.t. the full 3x3 product: only the upper triangle is read (strip_symmetric)
		G.m[0] = gcv[0]; G.m[1] = gcv[1]; G.m[2] = gcv[2]; G.m[3] = 0.f; G.m[4] = gcv[3];
		G.m[5] = gcv[4]; G.m[6] = 0.f; G.m[7] = 0.f; G.m[8] = gcv[5];
		Mat3 Gs;     // G + G^T
#pragma unroll
		for (int i = 0; i < 3; i++)
#pragma unroll
			for (int j = 0; j < 3; j++) Gs.m[3 * i + j] = G.m[3 * i + j] + G.m[3 * j + i];
		const Mat3 TS = mat3_mul(f.Tr, S);
		dTr = mat3_mul(Gs, TS);                              // d(T S T^T)/dT, S symmetric
		const Mat3 dS = mat3_mul(mat3_mul_tn(f.Tr, G), f.Tr); // T^T G T
		Mat3 dSs;
#pragma unroll
		for (int i = 0; i < 3; i++)
#pragma unroll
			for (int j = 0; j < 3; j++) dSs.m[3 * i + j] = dS.m[3 * i + j] + dS.m[3 * j + i];
		const Mat3 dL = mat3_mul(dSs, L);
		Mat3 dR;
#pragma unroll
		for (int k = 0; k < 3; k++) {
			const float ds = Rq.m[k] * dL.m[k] + Rq.m[3 + k] * dL.m[3 + k] + Rq.m[6 + k] * dL.m[6 + k];
			d_scaling[3 * (size_t)idx + k] = a.scale_modifier * ds;
#pragma unroll
			for (int i = 0; i < 3; i++) dR.m[3 * i + k] = dL.m[3 * i + k] * sc[k];
		}
		const float r = qr[0] * inv_norm, x = qr[1] * inv_norm, y = qr[2] * inv_norm, z = qr[3] * inv_norm;
		const float dr = 2.f * (z * (dR.m[3] - dR.m[1]) + y * (dR.m[2] - dR.m[6]) + x * (dR.m[7] - dR.m[5]));
		const float dx = 2.f * (y * (dR.m[1] + dR.m[3]) + z * (dR.m[2] + dR.m[6]) + r * (dR.m[7] - dR.m[5])) - 4.f * x * (dR.m[4] + dR.m[8]);
		const float dy = 2.f * (x * (dR.m[1] + dR.m[3]) + r * (dR.m[2] - dR.m[6]) + z * (dR.m[5] + dR.m[7])) - 4.f * y * (dR.m[0] + dR.m[8]);
		const float dz = 2.f * (r * (dR.m[3] - dR.m[1]) + x * (dR.m[2] + dR.m[6]) + y * (dR.m[5] + dR.m[7])) - 4.f * z * (dR.m[0] + dR.m[4]);
		const float dot = r * dr + x * dx + y * dy + z * dz;
		float4 o;
		o.x = (dr - r * dot) * inv_norm; o.y = (dx - x * dot) * inv_norm;
		o.z = (dy - y * dot) * inv_norm; o.w = (dz - z * dot) * inv_norm;
		*reinterpret_cast<float4*>(d_rotation + 4 * (size_t)idx) = o;
	} else {
		if (d_scaling) { d_scaling[3 * (size_t)idx] = 0.f; d_scaling[3 * (size_t)idx + 1] = 0.f; d_scaling[3 * (size_t)idx + 2] = 0.f; }
		if (d_rotation) *reinterpret_cast<float4*>(d_rotation + 4 * (size_t)idx) = make_float4(0.f, 0.f, 0.f, 0.f);
	}

	// ---- skinning chain --------------------------------------------------------------------------------
	float gs[3];
#pragma unroll
	for (int j = 0; j < 3; j++) gs[j] = sh.Rgi[3 * j] * gw[0] + sh.Rgi[3 * j + 1] * gw[1] + sh.Rgi[3 * j + 2] * gw[2];
	Mat3 Rg;
#pragma unroll
	for (int i = 0; i < 9; i++) Rg.m[i] = sh.Rg[i];
	const Mat3 gTr0 = mat3_mul_tn(Rg, dTr);
	Mat3 gRt = mat3_mul_nt(gTr0, f.Rbi);
#pragma unroll
	for (int i = 0; i < 3; i++)
#pragma unroll
		for (int j = 0; j < 3; j++) gRt.m[3 * i + j] += gs[i] * f.q[j];
	float gq[3];
	mat3_tvec(f.Rt, gs, gq);
	Mat3 gRbi = mat3_mul_tn(f.Rt, gTr0);
#pragma unroll
	for (int i = 0; i < 3; i++)
#pragma unroll
		for (int j = 0; j < 3; j++) gRbi.m[3 * i + j] += gq[i] * f.xb[j];
	float gx[3];
	mat3_tvec(f.Rbi, gq, gx);
#pragma unroll
	for (int i = 0; i < 3; i++) d_xyz[3 * (size_t)idx + i] = gx[i];
	if (d_poseoff_tgt) {
#pragma unroll
		for (int i = 0; i < 3; i++) atomicAdd(d_poseoff_tgt + 3 * (size_t)vid + i, gq[i]);
	}
	if (d_lbs_delta || d_A_tgt) {
		// gradient w.r.t. the blended 3x4 transforms, laid out like the table rows (r00 r01 r02 t0 | ...)
		const Mat3 tmp = mat3_mul_tn(f.Rbi, gRbi);
		const Mat3 gRb = mat3_mul_nt(tmp, f.Rbi);     // Rbi^T gRbi Rbi^T, sign applied below
		float gAb[12], gAt[12];
#pragma unroll
		for (int i = 0; i < 3; i++) {
#pragma unroll
			for (int j = 0; j < 3; j++) { gAb[4 * i + j] = -gRb.m[3 * i + j]; gAt[4 * i + j] = gRt.m[3 * i + j]; }
			gAb[4 * i + 3] = -gx[i];
			gAt[4 * i + 3] = gs[i];
		}
		if (d_A_tgt) {
#pragma unroll
			for (int e = 0; e < 12; e++) gAt_out[e] = gAt[e];
		}
		if (d_lbs_delta) {
			float gb[NJ];
			float mean = 0.f;
#pragma unroll
			for (int j = 0; j < NJ; j++) {
				float acc = 0.f;
#pragma unroll
				for (int e = 0; e < 12; e++) acc += gAb[e] * sh.Ab[j][e] + gAt[e] * sh.At[j][e];
				gb[j] = acc;
				mean += bw[j] * acc;
			}
			float4* o = reinterpret_cast<float4*>(d_lbs_delta + (size_t)idx * NJ);
#pragma unroll
			for (int k = 0; k < NJ / 4; k++)
				o[k] = make_float4(bw[4 * k] * (gb[4 * k] - mean), bw[4 * k + 1] * (gb[4 * k + 1] - mean),
				                   bw[4 * k + 2] * (gb[4 * k + 2] - mean), bw[4 * k + 3] * (gb[4 * k + 3] - mean));
		}
	}
}

// dL/dA_tgt[j][e] += sum over a warp's Gaussians of bw[j] * gAt[e]: a [24 x 32] . [32 x 12] product per warp.  The 32 weight
// rows and gradient rows are staged in shared memory (row stride 37 floats: conflict-free stores, broadcast loads) and
// every lane owns a 3 x 3 block of the 24 x 12 result (8 joint groups x 4 element groups): 32 x (6 loads + 9 FMA) per
// lane, then nine shared atomics per lane on addresses no other lane of the warp touches.  (Before: every thread issued
// 12 shared atomicAdd per joint it is bound to, the whole warp on the same 12 addresses -- a float atomicAdd on shared
// memory is a compare-and-swap loop, so a 32-way collision costs 32 rounds: the fused backward preprocess took 123 us at
// 20 k Gaussians with dL/dA_tgt wanted against 25 us without.)  Call with the whole warp converged; lanes without a
// Gaussian pass contributes = false.
constexpr int SKIN_DA_ROW = NJ + 12 + 1;
constexpr int SKIN_DA_STAGE = 32 * SKIN_DA_ROW;          // floats of staging per warp
__device__ __forceinline__ void skin_accumulate_dA(float* __restrict__ stage, float* accA, const float bw[NJ], const float gAt[12],
	bool contributes)
{
	const unsigned full = 0xffffffffu;
	if (!__any_sync(full, contributes)) return;
	const int lane = threadIdx.x & 31;
	float* row = stage + lane * SKIN_DA_ROW;
#pragma unroll
	for (int j = 0; j < NJ; j++) row[j] = contributes ? bw[j] : 0.f;
#pragma unroll
	for (int e = 0; e < 12; e++) row[NJ + e] = contributes ? gAt[e] : 0.f;
	__syncwarp();
	const int jg = 3 * (lane >> 2), eg = NJ + 3 * (lane & 3);
	float acc[9];
#pragma unroll
	for (int k = 0; k < 9; k++) acc[k] = 0.f;
#pragma unroll 8
	for (int i = 0; i < 32; i++) {
		const float* r = stage + i * SKIN_DA_ROW;
		const float w0 = r[jg], w1 = r[jg + 1], w2 = r[jg + 2], g0 = r[eg], g1 = r[eg + 1], g2 = r[eg + 2];
		acc[0] += w0 * g0; acc[1] += w0 * g1; acc[2] += w0 * g2;
		acc[3] += w1 * g0; acc[4] += w1 * g1; acc[5] += w1 * g2;
		acc[6] += w2 * g0; acc[7] += w2 * g1; acc[8] += w2 * g2;
	}
	__syncwarp();
#pragma unroll
	for (int a = 0; a < 3; a++)
#pragma unroll
		for (int b = 0; b < 3; b++)
			if (acc[3 * a + b] != 0.f) atomicAdd(&accA[(jg + a) * 12 + (eg - NJ) + b], acc[3 * a + b]);
}

__global__ void __launch_bounds__(SKIN_THREADS)
skin_backward_kernel(SkinArgs a, const float* __restrict__ g_means3D, const float* __restrict__ g_cov3D,
                     float* __restrict__ d_xyz, float* __restrict__ d_scaling, float* __restrict__ d_rotation,
                     float* __restrict__ d_lbs_delta, float* __restrict__ d_A_tgt, float* __restrict__ d_poseoff_tgt)
{
	__shared__ SkinShared sh;
	__shared__ float accA[NJ * 12];
	__shared__ float dA_stage[(SKIN_THREADS / 32) * SKIN_DA_STAGE];
	skin_stage(sh, a);
	if (d_A_tgt) {
		for (int i = threadIdx.x; i < NJ * 12; i += blockDim.x) accA[i] = 0.f;
		__syncthreads();
	}
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	float bw[NJ], gAt[12];
	if (idx < a.P) {
		const int vid = __ldg(a.vert_ids + idx);
		const SkinOwn own = skin_own(a, idx);
		float gw[3], gcv[6];
#pragma unroll
		for (int i = 0; i < 3; i++) gw[i] = __ldg(g_means3D + 3 * (size_t)idx + i);
#pragma unroll
		for (int i = 0; i < 6; i++) gcv[i] = g_cov3D ? __ldg(g_cov3D + 6 * (size_t)idx + i) : 0.f;
		skin_weights(a, idx, vid, bw);
		SkinFwd f;
		skin_point(sh, a, own.x, vid, bw, f);
		skin_backward_point(sh, a, idx, vid, own, bw, f, gw, gcv, g_cov3D != nullptr, d_xyz, d_scaling, d_rotation, d_lbs_delta,
			gAt, d_A_tgt != nullptr, d_poseoff_tgt);
	}
	if (d_A_tgt) {
		skin_accumulate_dA(dA_stage + (threadIdx.x >> 5) * SKIN_DA_STAGE, accA, bw, gAt, idx < a.P);
		__syncthreads();
		for (int i = threadIdx.x; i < NJ * 12; i += blockDim.x) {
			const float v = accA[i];
			if (v != 0.f) atomicAdd(d_A_tgt + (i / 12) * 16 + (i % 12), v);
		}
	}
}

} // namespace moss
