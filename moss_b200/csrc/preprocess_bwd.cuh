// K7: per-Gaussian backward preprocess -- one fused kernel for what the reference runs as two
// (computeCov2DCUDA, DGR/cuda_rasterizer/backward.cu:144-274, then preprocessCUDA, :346-396, with the SH and
// cov3D chain rules :20-139 and :278-341).  Every output row is written (zeros for culled Gaussians and for
// SH coefficients above the active degree), so the caller never zero-fills dL_dmean3D / dL_dcov3D / dL_dsh /
// dL_dscale / dL_drot (the reference memsets 9 tensors per call, DGR/rasterize_points.cu:158-166).
// Gradients are toleranced (1e-3 relative), so plain fp32 expressions are used here.
#pragma once
#include "common.cuh"
#include "preprocess.cuh"
#include "push_exchange.cuh"

namespace moss {

__device__ __forceinline__ void store_quat_routed(const PushCfg& pc, float* out, int i, float4 v)
{
	if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) { *reinterpret_cast<float4*>(route(pc, out + 4 * i)) = v; return; }
	*route(pc, out + 4 * i) = v.x; *route(pc, out + 4 * i + 1) = v.y; *route(pc, out + 4 * i + 2) = v.z; *route(pc, out + 4 * i + 3) = v.w;
}

// Where the per-Gaussian parameter gradients of view v go.  One view: the caller's tensors.  A batch of V views: row
// v of a scratch area (ptr + v * view_stride), summed over the views afterwards by sum_views_kernel in a fixed order.
struct ParamGradOut {
	float* color;     // [P,3]  (colors_precomp path)
	float* opacity;   // [P]
	float* mean3D;    // [P,3]
	float* cov3D;     // [P,6]
	float* sh;        // [P,M,3] or null
	float* scale;     // [P,3] or null
	float* rot;       // [P,4] or null
	size_t view_stride;   // floats between consecutive views' rows (0 for a single view)
};

// Where the fused skin + preprocess backward (SKIN = true) puts the front end's gradients: rows per Gaussian, and the two
// accumulators of the per-frame tables (cleared by the caller).
struct SkinGradOut {
	float* d_xyz;          // [P,3]
	float* d_scaling;      // [P,3]
	float* d_rotation;     // [P,4]
	float* d_lbs_delta;    // [P,24] or null
	float* d_A_tgt;        // [24,16] or null
	float* d_poseoff_tgt;  // [V,3] or null
};

// grid = (ceil(P / PRE_THREADS), V).  SKIN: single view, no exchange; means3D / cov3D are recomputed from the skinning
// inputs and dL/dmeans3D, dL/dcov3D go straight into the skinning backward (skin.cuh) without touching HBM.
// Register caps: none below the architectural 255 pays.  Plain kernel: 80-96 spill and lose 6-13 % at Stress
// (profiles/r02_h_ab_regcaps_preprocess_sort.txt).  Fused (SKIN) variant: 233 registers = two resident blocks; a cap
// of 168 buys a third, spills 136 bytes and trades 100 k Gaussians 54.8 -> 49.7 us for 20 k 25.3 -> 27.2 us.
#ifndef MOSS_PRE_BWD_SKIN_MAXREG
#define MOSS_PRE_BWD_SKIN_MAXREG 255
#endif
#define MOSS_PRE_BWD_REGCAP __maxnreg__(SKIN ? MOSS_PRE_BWD_SKIN_MAXREG : 255)
template <bool SKIN>
__global__ void MOSS_PRE_BWD_REGCAP
preprocess_backward_kernel(const __grid_constant__ ViewBatch vb,
	const float* __restrict__ means3D, const int* __restrict__ radii, const float* __restrict__ shs,
	const float* __restrict__ scales, const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
	GeomWS g,
	float* __restrict__ dL_dmean2D_all, float* __restrict__ dL_dconic_all, ParamGradOut po,
	int deterministic, const uint32_t* __restrict__ counters, const __grid_constant__ PushCfg pc,
	const SkinFused sk, const SkinGradOut sg)
{
	// All global reads of the Gaussian are issued before the first data-dependent branch (one DRAM round trip instead
	// of a chain; see preprocess_forward_kernel) -- except the accumulator row, which only visible Gaussians own.
	const int P = vb.P;
	const int view = blockIdx.y;
	const ViewCam& cam = vb.cam[view];
	const int idx = blockIdx.x * PRE_THREADS + threadIdx.x;
	const bool active = idx < P;
	const int li = active ? idx : 0;
	const size_t vid = (size_t)view * P + li;
	const int M = vb.M, D = vb.D;
	const int radius = radii[vid];
	const size_t vo = (size_t)view * po.view_stride;
	float* __restrict__ dL_dmean2D = dL_dmean2D_all + 3 * (size_t)view * P;
	float* __restrict__ dL_dconic = dL_dconic_all ? dL_dconic_all + 4 * (size_t)view * P : nullptr;
	float* __restrict__ dL_dcolor = po.color + vo;
	float* __restrict__ dL_dopacity = po.opacity + vo;
	float* __restrict__ dL_dmean3D = SKIN ? nullptr : po.mean3D + vo;
	float* __restrict__ dL_dcov3D = SKIN ? nullptr : po.cov3D + vo;
	float* __restrict__ dL_dsh = po.sh ? po.sh + vo : nullptr;
	float* __restrict__ dL_dscale = po.scale ? po.scale + vo : nullptr;
	float* __restrict__ dL_drot = po.rot ? po.rot + vo : nullptr;
	float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0;
	float acc2 = 0.f;
	if (radius > 0) {
		if (!deterministic) {
			acc0 = g.grad_acc[GRAD_ACC_VEC * vid]; acc1 = g.grad_acc[GRAD_ACC_VEC * vid + 1];
			acc2 = g.grad_acc[GRAD_ACC_VEC * vid + 2].x;
		} else {
			// nine int64 fixed-point sums (blend.cuh) -> float
			const float inv = det_scale_from_bits(counters[6], true);
			const long long* a = reinterpret_cast<const long long*>(g.grad_acc + GRAD_ACC_VEC * vid);
			const long long* e = reinterpret_cast<const long long*>(g.grad_ext + GRAD_EXT_VEC * vid);
			acc0 = make_float4((float)a[0] * inv, (float)a[1] * inv, (float)a[2] * inv, (float)a[3] * inv);
			acc1 = make_float4((float)a[4] * inv, (float)a[5] * inv, (float)e[0] * inv, (float)e[1] * inv);
			acc2 = (float)e[2] * inv;
		}
	}
	float mx = 0.f, my = 0.f, mz = 0.f, V0 = 0.f, V1 = 0.f, V2 = 0.f, V3 = 0.f, V4 = 0.f, V5 = 0.f;
	int sk_vid = 0;
	SkinOwn sk_own;
	if constexpr (SKIN) {
		sk_vid = __ldg(sk.a.vert_ids + li);
		sk_own = skin_own(sk.a, li);
	} else {
		mx = means3D[3 * li]; my = means3D[3 * li + 1]; mz = means3D[3 * li + 2];
		const float* c3 = cov3D_precomp ? (cov3D_precomp + 6 * (size_t)li) : (g.cov3D + 6 * vid);
		V0 = c3[0]; V1 = c3[1]; V2 = c3[2]; V3 = c3[3]; V4 = c3[4]; V5 = c3[5];
	}
	const uint8_t cl = g.clamped[vid];
	const bool sh_vec = shs && ((reinterpret_cast<uintptr_t>(shs) & 15) == 0) && ((3 * M) % 4 == 0);
	const bool dsh_vec = shs && M == 16 && ((reinterpret_cast<uintptr_t>(dL_dsh) & 15) == 0);
	float sh[48];
	if (shs) sh_load(D, M, shs + (size_t)li * 3 * M, sh_vec, sh);
	float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
	float sc_in[3] = {0.f, 0.f, 0.f};
	if (!SKIN && scales) {
		q = load_quat(rotations, li);
		sc_in[0] = scales[3 * li]; sc_in[1] = scales[3 * li + 1]; sc_in[2] = scales[3 * li + 2];
	}

	__shared__ float v[16], pr[16], s_cam[3];
	if (threadIdx.x < 16) v[threadIdx.x] = cam.view[threadIdx.x];
	else if (threadIdx.x < 32) pr[threadIdx.x - 16] = cam.proj[threadIdx.x - 16];
	else if (threadIdx.x < 35 && shs) s_cam[threadIdx.x - 32] = cam.campos[threadIdx.x - 32];
	__syncthreads();
	__shared__ SkinShared sk_sh;         // SKIN only (unused allocations are dropped by the compiler otherwise)
	__shared__ float accA[NJ * 12];
	__shared__ float dA_stage[SKIN ? (PRE_THREADS / 32) * SKIN_DA_STAGE : 1];
	if constexpr (SKIN) {
		skin_stage(sk_sh, sk.a);
		if (sg.d_A_tgt) {
			for (int i = threadIdx.x; i < NJ * 12; i += blockDim.x) accA[i] = 0.f;
			__syncthreads();
		}
	}
	// The plain kernel leaves through early returns.  In the fused form every thread has to come back for the block-level
	// flush of dL/dA_tgt, so there the exits clear `live` instead.  (No lambda around the body: capturing the locals by
	// reference cost the plain kernel 11 registers and ~8 % at Stress.)
	if constexpr (!SKIN) { if (!active) return; }
	bool live = active;
	float sk_bw[NJ], sk_gAt[12];
	SkinFwd sk_f;
	if (live) {
	if constexpr (SKIN) {
		// recompute the front end's forward (nothing of it was stored): world position and covariance
		skin_weights(sk.a, li, sk_vid, sk_bw);
		skin_point(sk_sh, sk.a, sk_own.x, sk_vid, sk_bw, sk_f);
		float world[3], c6[6];
		skin_world(sk_sh, sk_f, world);
		skin_cov(sk_own, sk_f, c6);
		mx = world[0]; my = world[1]; mz = world[2];
		V0 = c6[0]; V1 = c6[1]; V2 = c6[2]; V3 = c6[3]; V4 = c6[4]; V5 = c6[5];
	}

	// The blend backward accumulated into g.grad_acc (vector atomics); unpack it into the caller's tensors, which are
	// therefore fully written here (no zero-fill beforehand), and leave the accumulators zero for the next backward.
	if (radius > 0) {
#pragma unroll
		for (int k = 0; k < GRAD_ACC_VEC; k++) g.grad_acc[GRAD_ACC_VEC * vid + k] = make_float4(0.f, 0.f, 0.f, 0.f);
		if (deterministic) {
#pragma unroll
			for (int k = 0; k < GRAD_EXT_VEC; k++) g.grad_ext[GRAD_EXT_VEC * vid + k] = make_float4(0.f, 0.f, 0.f, 0.f);
		}
	}
	// Every store of a parameter-gradient element goes through route(): with a data-parallel exchange attached the
	// element lands in the staging area of the rank that owns it (push_exchange.cuh), otherwise in the caller's tensor.
	dL_dcolor[3 * idx] = acc0.x; dL_dcolor[3 * idx + 1] = acc0.y; dL_dcolor[3 * idx + 2] = acc0.z;
	*route(pc, dL_dopacity + idx) = acc0.w;
	*route(pc, dL_dmean2D + 3 * idx) = acc1.x; *route(pc, dL_dmean2D + 3 * idx + 1) = acc1.y; *route(pc, dL_dmean2D + 3 * idx + 2) = 0.f;
	if (dL_dconic) { dL_dconic[4 * idx] = acc1.z; dL_dconic[4 * idx + 1] = acc1.w; dL_dconic[4 * idx + 2] = 0.f; dL_dconic[4 * idx + 3] = acc2; }

	if (!(radius > 0)) {
		if constexpr (SKIN) {
			// a culled Gaussian: every front-end gradient row is zero
#pragma unroll
			for (int k = 0; k < 3; k++) { sg.d_xyz[3 * (size_t)idx + k] = 0.f; sg.d_scaling[3 * (size_t)idx + k] = 0.f; }
			*reinterpret_cast<float4*>(sg.d_rotation + 4 * (size_t)idx) = make_float4(0.f, 0.f, 0.f, 0.f);
			if (sg.d_lbs_delta) {
#pragma unroll
				for (int k = 0; k < NJ / 4; k++) reinterpret_cast<float4*>(sg.d_lbs_delta + (size_t)idx * NJ)[k] = make_float4(0.f, 0.f, 0.f, 0.f);
			}
		} else {
#pragma unroll
			for (int k = 0; k < 3; k++) *route(pc, dL_dmean3D + 3 * idx + k) = 0.f;
#pragma unroll
			for (int k = 0; k < 6; k++) *route(pc, dL_dcov3D + 6 * idx + k) = 0.f;
		}
		if (shs) {
			if (dsh_vec) {
				float* d = dL_dsh + (size_t)idx * 48;
#pragma unroll
				for (int k = 0; k < 12; k++) *reinterpret_cast<float4*>(route(pc, d + 4 * k)) = make_float4(0.f, 0.f, 0.f, 0.f);
			} else {
				for (int k = 0; k < 3 * M; k++) *route(pc, dL_dsh + (size_t)idx * 3 * M + k) = 0.f;
			}
		}
		if (!SKIN && scales) {
#pragma unroll
			for (int k = 0; k < 3; k++) *route(pc, dL_dscale + 3 * idx + k) = 0.f;
			store_quat_routed(pc, dL_drot, idx, make_float4(0.f, 0.f, 0.f, 0.f));
		}
		if (pc.numel) __threadfence_system();
		if constexpr (SKIN) live = false; else return;
	}
	}
	if (live) {
	// ---- conic -> cov2D -> cov3D and mean (through J) ----------------------------------------------
	const float dcx = acc1.z, dcy = acc1.w, dcz = acc2;
	float tx = v[0] * mx + v[4] * my + v[8] * mz + v[12];
	float ty = v[1] * mx + v[5] * my + v[9] * mz + v[13];
	const float tz = v[2] * mx + v[6] * my + v[10] * mz + v[14];
	const float limx = 1.3f * cam.tanfovx, limy = 1.3f * cam.tanfovy;
	const float txtz = tx / tz, tytz = ty / tz;
	tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
	ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
	const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
	const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
	const float hx = cam.focal_x, hy = cam.focal_y;

	const float J00 = hx / tz, J02 = -(hx * tx) / (tz * tz);
	const float J11 = hy / tz, J12 = -(hy * ty) / (tz * tz);
	// T = W*J : two non-zero columns (T0*, T1*)
	const float T00 = v[0] * J00 + v[2] * J02, T01 = v[4] * J00 + v[6] * J02, T02 = v[8] * J00 + v[10] * J02;
	const float T10 = v[1] * J11 + v[2] * J12, T11 = v[5] * J11 + v[6] * J12, T12 = v[9] * J11 + v[10] * J12;
	// V*T columns
	const float VT00 = T00 * V0 + T01 * V1 + T02 * V2;
	const float VT01 = T00 * V1 + T01 * V3 + T02 * V4;
	const float VT02 = T00 * V2 + T01 * V4 + T02 * V5;
	const float VT10 = T10 * V0 + T11 * V1 + T12 * V2;
	const float VT11 = T10 * V1 + T11 * V3 + T12 * V4;
	const float VT12 = T10 * V2 + T11 * V4 + T12 * V5;
	const float a = (T00 * VT00 + T01 * VT01 + T02 * VT02) + 0.3f;
	const float bq = T00 * VT10 + T01 * VT11 + T02 * VT12;
	const float c = (T10 * VT10 + T11 * VT11 + T12 * VT12) + 0.3f;

	const float denom = a * c - bq * bq;
	float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
	const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
	float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
	if (denom2inv != 0.f) {
		dL_da = denom2inv * (-c * c * dcx + 2 * bq * c * dcy + (denom - a * c) * dcz);
		dL_dc = denom2inv * (-a * a * dcz + 2 * a * bq * dcy + (denom - a * c) * dcx);
		dL_db = denom2inv * 2 * (bq * c * dcx - (denom + 2 * bq * bq) * dcy + a * bq * dcz);
		dcov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
		dcov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
		dcov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
		dcov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
		dcov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
		dcov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
	}
	if constexpr (!SKIN) {
#pragma unroll
		for (int k = 0; k < 6; k++) *route(pc, dL_dcov3D + 6 * idx + k) = dcov[k];
	}

	const float dT00 = 2 * VT00 * dL_da + VT10 * dL_db;
	const float dT01 = 2 * VT01 * dL_da + VT11 * dL_db;
	const float dT02 = 2 * VT02 * dL_da + VT12 * dL_db;
	const float dT10 = 2 * VT10 * dL_dc + VT00 * dL_db;
	const float dT11 = 2 * VT11 * dL_dc + VT01 * dL_db;
	const float dT12 = 2 * VT12 * dL_dc + VT02 * dL_db;
	const float dJ00 = v[0] * dT00 + v[4] * dT01 + v[8] * dT02;
	const float dJ02 = v[2] * dT00 + v[6] * dT01 + v[10] * dT02;
	const float dJ11 = v[1] * dT10 + v[5] * dT11 + v[9] * dT12;
	const float dJ12 = v[2] * dT10 + v[6] * dT11 + v[10] * dT12;
	const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
	const float dtx = x_grad_mul * -hx * itz2 * dJ02;
	const float dty = y_grad_mul * -hy * itz2 * dJ12;
	// z-dependence of the frustum clamp is ignored exactly like the reference (backward.cu:264)
	const float dtz = -hx * itz2 * dJ00 - hy * itz2 * dJ11 + (2 * hx * tx) * itz3 * dJ02 + (2 * hy * ty) * itz3 * dJ12;
	float gmx = v[0] * dtx + v[1] * dty + v[2] * dtz;
	float gmy = v[4] * dtx + v[5] * dty + v[6] * dtz;
	float gmz = v[8] * dtx + v[9] * dty + v[10] * dtz;

	// ---- 2D mean -> 3D mean through the projection ---------------------------------------------------
	{
		const float m_w = 1.0f / ((pr[3] * mx + pr[7] * my + pr[11] * mz + pr[15]) + 0.0000001f);
		const float mul1 = (pr[0] * mx + pr[4] * my + pr[8] * mz + pr[12]) * m_w * m_w;
		const float mul2 = (pr[1] * mx + pr[5] * my + pr[9] * mz + pr[13]) * m_w * m_w;
		const float d2x = acc1.x, d2y = acc1.y;
		gmx += (pr[0] * m_w - pr[3] * mul1) * d2x + (pr[1] * m_w - pr[3] * mul2) * d2y;
		gmy += (pr[4] * m_w - pr[7] * mul1) * d2x + (pr[5] * m_w - pr[7] * mul2) * d2y;
		gmz += (pr[8] * m_w - pr[11] * mul1) * d2x + (pr[9] * m_w - pr[11] * mul2) * d2y;
	}

	// ---- colour -> SH coefficients and view direction -> mean ---------------------------------------
	if (shs) {
		float* dsh = dL_dsh + (size_t)idx * 3 * M;
		const float ox = mx - s_cam[0], oy = my - s_cam[1], oz = mz - s_cam[2];
		const float ilen = 1.f / sqrtf(ox * ox + oy * oy + oz * oz);
		const float x = ox * ilen, y = oy * ilen, z = oz * ilen;
		float dRGB[3];
#pragma unroll
		for (int ch = 0; ch < 3; ch++) dRGB[ch] = ((cl >> ch) & 1) ? 0.f : (ch == 0 ? acc0.x : (ch == 1 ? acc0.y : acc0.z));

		float basis[16];
#pragma unroll
		for (int k = 0; k < 16; k++) basis[k] = 0.f;
		float dx3[3] = {0.f, 0.f, 0.f}, dy3[3] = {0.f, 0.f, 0.f}, dz3[3] = {0.f, 0.f, 0.f};
		basis[0] = MOSS_SH_C0;
		if (D > 0) {
			basis[1] = -MOSS_SH_C1 * y; basis[2] = MOSS_SH_C1 * z; basis[3] = -MOSS_SH_C1 * x;
#pragma unroll
			for (int ch = 0; ch < 3; ch++) {
				dx3[ch] = -MOSS_SH_C1 * sh[9 + ch];
				dy3[ch] = -MOSS_SH_C1 * sh[3 + ch];
				dz3[ch] = MOSS_SH_C1 * sh[6 + ch];
			}
			if (D > 1) {
				const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
				basis[4] = MOSS_SH_C2_0 * xy; basis[5] = MOSS_SH_C2_1 * yz;
				basis[6] = MOSS_SH_C2_2 * (2.f * zz - xx - yy);
				basis[7] = MOSS_SH_C2_3 * xz; basis[8] = MOSS_SH_C2_4 * (xx - yy);
#pragma unroll
				for (int ch = 0; ch < 3; ch++) {
					const float s4 = sh[12 + ch], s5 = sh[15 + ch], s6 = sh[18 + ch], s7 = sh[21 + ch], s8 = sh[24 + ch];
					dx3[ch] += MOSS_SH_C2_0 * y * s4 + MOSS_SH_C2_2 * 2.f * -x * s6 + MOSS_SH_C2_3 * z * s7 + MOSS_SH_C2_4 * 2.f * x * s8;
					dy3[ch] += MOSS_SH_C2_0 * x * s4 + MOSS_SH_C2_1 * z * s5 + MOSS_SH_C2_2 * 2.f * -y * s6 + MOSS_SH_C2_4 * 2.f * -y * s8;
					dz3[ch] += MOSS_SH_C2_1 * y * s5 + MOSS_SH_C2_2 * 2.f * 2.f * z * s6 + MOSS_SH_C2_3 * x * s7;
				}
				if (D > 2) {
					basis[9] = MOSS_SH_C3_0 * y * (3.f * xx - yy);
					basis[10] = MOSS_SH_C3_1 * xy * z;
					basis[11] = MOSS_SH_C3_2 * y * (4.f * zz - xx - yy);
					basis[12] = MOSS_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
					basis[13] = MOSS_SH_C3_4 * x * (4.f * zz - xx - yy);
					basis[14] = MOSS_SH_C3_5 * z * (xx - yy);
					basis[15] = MOSS_SH_C3_6 * x * (xx - 3.f * yy);
#pragma unroll
					for (int ch = 0; ch < 3; ch++) {
						const float s9 = sh[27 + ch], s10 = sh[30 + ch], s11 = sh[33 + ch], s12 = sh[36 + ch],
							s13 = sh[39 + ch], s14 = sh[42 + ch], s15 = sh[45 + ch];
						dx3[ch] += MOSS_SH_C3_0 * s9 * 3.f * 2.f * xy + MOSS_SH_C3_1 * s10 * yz + MOSS_SH_C3_2 * s11 * -2.f * xy +
							MOSS_SH_C3_3 * s12 * -3.f * 2.f * xz + MOSS_SH_C3_4 * s13 * (-3.f * xx + 4.f * zz - yy) +
							MOSS_SH_C3_5 * s14 * 2.f * xz + MOSS_SH_C3_6 * s15 * 3.f * (xx - yy);
						dy3[ch] += MOSS_SH_C3_0 * s9 * 3.f * (xx - yy) + MOSS_SH_C3_1 * s10 * xz +
							MOSS_SH_C3_2 * s11 * (-3.f * yy + 4.f * zz - xx) + MOSS_SH_C3_3 * s12 * -3.f * 2.f * yz +
							MOSS_SH_C3_4 * s13 * -2.f * xy + MOSS_SH_C3_5 * s14 * -2.f * yz + MOSS_SH_C3_6 * s15 * -3.f * 2.f * xy;
						dz3[ch] += MOSS_SH_C3_1 * s10 * xy + MOSS_SH_C3_2 * s11 * 4.f * 2.f * yz +
							MOSS_SH_C3_3 * s12 * 3.f * (2.f * zz - xx - yy) + MOSS_SH_C3_4 * s13 * 4.f * 2.f * xz +
							MOSS_SH_C3_5 * s14 * (xx - yy);
					}
				}
			}
		}
		const int ncoef = (D + 1) * (D + 1);
		if (dsh_vec) {
			// the 192 B row as 12 vector stores (48 scalar stores cost 4x the LSU time: every lane owns its sectors)
			float o[48];
#pragma unroll
			for (int k = 0; k < 16; k++) {
				const float bk = (k < ncoef) ? basis[k] : 0.f;
				o[3 * k] = bk * dRGB[0]; o[3 * k + 1] = bk * dRGB[1]; o[3 * k + 2] = bk * dRGB[2];
			}
#pragma unroll
			for (int k = 0; k < 12; k++)
				*reinterpret_cast<float4*>(route(pc, dsh + 4 * k)) = make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
		} else {
#pragma unroll
			for (int k = 0; k < 16; k++) {
				if (k < M) {
					const float bk = (k < ncoef) ? basis[k] : 0.f;
					*route(pc, dsh + 3 * k) = bk * dRGB[0]; *route(pc, dsh + 3 * k + 1) = bk * dRGB[1]; *route(pc, dsh + 3 * k + 2) = bk * dRGB[2];
				}
			}
			for (int k = 16; k < M; k++) { *route(pc, dsh + 3 * k) = 0.f; *route(pc, dsh + 3 * k + 1) = 0.f; *route(pc, dsh + 3 * k + 2) = 0.f; }
		}

		const float ddx = dx3[0] * dRGB[0] + dx3[1] * dRGB[1] + dx3[2] * dRGB[2];
		const float ddy = dy3[0] * dRGB[0] + dy3[1] * dRGB[1] + dy3[2] * dRGB[2];
		const float ddz = dz3[0] * dRGB[0] + dz3[1] * dRGB[1] + dz3[2] * dRGB[2];
		// d normalize(v)/dv applied to dL_ddir (DGR auxiliary.h:107-117)
		const float sum2 = ox * ox + oy * oy + oz * oz;
		const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
		gmx += ((sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
		gmy += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
		gmz += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
	}
	if constexpr (SKIN) {
		// dL/dmeans3D and dL/dcov3D never leave the registers: straight into the skinning backward
		const float gw[3] = {gmx, gmy, gmz};
		skin_backward_point(sk_sh, sk.a, idx, sk_vid, sk_own, sk_bw, sk_f, gw, dcov, true, sg.d_xyz, sg.d_scaling, sg.d_rotation,
			sg.d_lbs_delta, sk_gAt, sg.d_A_tgt != nullptr, sg.d_poseoff_tgt);
	} else {
		*route(pc, dL_dmean3D + 3 * idx) = gmx; *route(pc, dL_dmean3D + 3 * idx + 1) = gmy; *route(pc, dL_dmean3D + 3 * idx + 2) = gmz;
	}

	// ---- cov3D -> scale, (un-normalised) quaternion --------------------------------------------------
	if (!SKIN && scales) {
		const float r = q.x, x = q.y, y = q.z, z = q.w;
		const float s[3] = {vb.scale_modifier * sc_in[0], vb.scale_modifier * sc_in[1], vb.scale_modifier * sc_in[2]};
		// R[i][j]: column i, row j (GLM listing of DGR backward.cu:288-292)
		const float R[3][3] = {
			{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
			{2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
			{2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
		const float dS[3][3] = {
			{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
			{0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
			{0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
		// dL_dM[i][j] = sum_k 2*M[k][j]*dS[i][k],  M[k][j] = s_j * R[k][j];  dMt[a][b] = dL_dM[b][a]
		float dMt[3][3];
#pragma unroll
		for (int i = 0; i < 3; i++)
#pragma unroll
			for (int j = 0; j < 3; j++)
				dMt[j][i] = 2.0f * s[j] * R[0][j] * dS[i][0] + 2.0f * s[j] * R[1][j] * dS[i][1] + 2.0f * s[j] * R[2][j] * dS[i][2];
		// dL_dscale_k = dot(Rt[k], dMt[k]),  Rt[k][j] = R[j][k]   (not multiplied by the modifier, as the reference)
#pragma unroll
		for (int k = 0; k < 3; k++)
			*route(pc, dL_dscale + 3 * idx + k) = R[0][k] * dMt[k][0] + R[1][k] * dMt[k][1] + R[2][k] * dMt[k][2];
#pragma unroll
		for (int k = 0; k < 3; k++) { dMt[k][0] *= s[k]; dMt[k][1] *= s[k]; dMt[k][2] *= s[k]; }
		float4 dq;
		dq.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
		dq.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
		dq.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
		dq.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
		store_quat_routed(pc, dL_drot, idx, dq);   // no normalisation Jacobian (DGR backward.cu:340)
	}
	if (pc.numel) __threadfence_system();   // pushed rows are on their way before this thread retires
	}
	if constexpr (SKIN) {
		if (sg.d_A_tgt) {
			skin_accumulate_dA(dA_stage + (threadIdx.x >> 5) * SKIN_DA_STAGE, accA, sk_bw, sk_gAt, live);
			__syncthreads();
			for (int i = threadIdx.x; i < NJ * 12; i += blockDim.x) {
				const float v = accA[i];
				if (v != 0.f) atomicAdd(sg.d_A_tgt + (i / 12) * 16 + (i % 12), v);
			}
		}
	}
}

// Batched backward, V > 1: out[j] = sum_v rows[v*stride + j] over the views, in view order (fixed order: the sum is
// reproducible).  One launch covers the (up to 7) parameter-gradient arrays: blockIdx.y picks the array.
struct SumViewsArgs {
	float* dst[7];
	const float* src[7];       // row of view 0
	size_t n[7];               // floats
	size_t view_stride;
	int V;
};
constexpr int SUMV_THREADS = 256;
__global__ void __launch_bounds__(SUMV_THREADS)
sum_views_kernel(const __grid_constant__ SumViewsArgs a, const __grid_constant__ PushCfg pc)
{
	const int k = blockIdx.y;
	float* __restrict__ dst = a.dst[k];
	const float* __restrict__ src = a.src[k];
	const size_t n = a.n[k];
	if (!dst || n == 0) return;
	const size_t stride = (size_t)gridDim.x * SUMV_THREADS;
	if ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0 && (a.view_stride & 3) == 0) {
		const size_t n4 = n >> 2;
		for (size_t i = (size_t)blockIdx.x * SUMV_THREADS + threadIdx.x; i < n4; i += stride) {
			float4 s = reinterpret_cast<const float4*>(src)[i];
			for (int v = 1; v < a.V; v++) {
				const float4 x = reinterpret_cast<const float4*>(src + (size_t)v * a.view_stride)[i];
				s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
			}
			*reinterpret_cast<float4*>(route(pc, dst + 4 * i)) = s;
		}
		for (size_t i = (n4 << 2) + (size_t)blockIdx.x * SUMV_THREADS + threadIdx.x; i < n; i += stride) {
			float s = src[i];
			for (int v = 1; v < a.V; v++) s += src[(size_t)v * a.view_stride + i];
			*route(pc, dst + i) = s;
		}
	} else {
		for (size_t i = (size_t)blockIdx.x * SUMV_THREADS + threadIdx.x; i < n; i += stride) {
			float s = src[i];
			for (int v = 1; v < a.V; v++) s += src[(size_t)v * a.view_stride + i];
			*route(pc, dst + i) = s;
		}
	}
	if (pc.numel) __threadfence_system();
}

} // namespace moss
