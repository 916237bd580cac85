// Per-frame SMPL tables as native kernels: what MOSS's render() recomputes for every
// frame before it touches a Gaussian --
//   batch_rodrigues                       scene/gaussian_model.py:1033-1064   24 axis-angle vectors -> rotation matrices
//   optional pose correction              :874-877                            rot[j] = rodrigues[j] @ correct_Rs[j-1], j >= 1
//   get_transform_params_torch            :998-1031   v_shaped = v_template + shapedirs . betas, joints = J_regressor . v_shaped
//   get_rigid_transformation_torch        :965-995    the 24-joint kinematic chain (a Python loop of 4x4 products),
//                                                     A[j] = chain[j] with the rest joint removed from its translation
//   pose blend shapes                     :849-856 / :872-888   pose_offsets = (rot[1:] - I).flatten() @ posedirs^T   (17 MB of posedirs)
//   shape blend shapes                    :862-866              shape_offset = shapedirs . betas
// -- about 60 small torch kernels, two Python loops over the joints and two 207 x 20 670 matvecs per frame in the
// reference.  Here: ONE launch forward, ONE backward.
//
// Forward, grid = shape blocks + pose blocks:
//   shape blocks (one thread per SMPL vertex): shape_offset row, v_shaped, the block's partial J_regressor sums; the
//     LAST shape block to finish (ticket) adds the partials in block order (fixed order -> reproducible), runs the
//     Rodrigues formulas and walks the chain with 24 threads -> A [24,4,4], rot_mats [24,3,3];
//   pose blocks (one warp per posedirs row, two rows in flight per warp): each block rebuilds the 207 pose features in
//     shared memory (24 Rodrigues evaluations -- cheaper than a dependent launch) and streams its rows: HBM-bound.
// Backward (gradients w.r.t. poses and correct_Rs from dL/dA and dL/dpose_offsets; shapes and the SMPL tables are
// constants of a run): pose blocks form the partial d(pose feature) sums of their rows (the transposed matvec), the last
// block adds them in block order, back-propagates the chain (children before parents) and the Rodrigues formula.
#pragma once
#include "common.cuh"
#include "skin.cuh"

namespace moss {

constexpr int SMPL_THREADS = 256;
constexpr int SMPL_ROWS_PER_BLOCK = 64;      // posedirs rows per pose block (8 warps x 8 rows)
constexpr int NPF = (NJ - 1) * 9;            // 207 pose features

struct SmplFrameArgs {
	int V;                      // SMPL vertices (6890)
	int n_shapes;               // betas used (<= 10)
	const float* v_template;    // [V,3]
	const float* shapedirs;     // [V,3,shape_stride]
	int shape_stride;           // 10
	const float* posedirs;      // [V*3, 207] (reference layout [V,3,207])
	const float* J_regressor;   // [24,V]
	const int*   parents;       // [24]  (kintree_table[0]; parents[0] ignored)
	const float* poses;         // [72] axis-angle
	const float* shapes;        // [n_shapes]
	const float* correct_Rs;    // [23,9] or null
	float* A;                   // [24,16]
	float* rot_mats;            // [24,9]  (after the pose correction) or null
	float* joints;              // [24,3]  J_regressor . v_shaped (the backward reads it) or null
	float* poseoff;             // [V,3] or null (skip the pose blocks' work)
	float* shapeoff;            // [V,3] or null
	float* partial;             // scratch: [shape_blocks][72] joint partial sums
	unsigned* ticket;           // scratch: zero on entry, zero on exit
	int shape_blocks;
};

// R = I + sin(a) K + (1 - cos(a)) K K with a = |r + 1e-8| and K built from r / a (:1047-1063; the 1e-8 sits inside the norm)
__device__ __forceinline__ Mat3 rodrigues(float rx, float ry, float rz)
{
	const float ex = rx + 1e-8f, ey = ry + 1e-8f, ez = rz + 1e-8f;
	const float angle = sqrtf(ex * ex + ey * ey + ez * ez);
	const float kx = rx / angle, ky = ry / angle, kz = rz / angle;
	float s, c;
	sincosf(angle, &s, &c);
	const Mat3 K = {{0.f, -kz, ky, kz, 0.f, -kx, -ky, kx, 0.f}};
	const Mat3 KK = mat3_mul(K, K);
	Mat3 R;
#pragma unroll
	for (int i = 0; i < 9; i++) R.m[i] = ((i % 4 == 0) ? 1.f : 0.f) + s * K.m[i] + (1.f - c) * KK.m[i];
	return R;
}

// rotation of joint j after the optional correction (:874-877): thread-local, used by both block roles
__device__ __forceinline__ Mat3 joint_rotation(const float* __restrict__ poses, const float* __restrict__ correct_Rs, int j)
{
	Mat3 R = rodrigues(poses[3 * j], poses[3 * j + 1], poses[3 * j + 2]);
	if (correct_Rs && j >= 1) {
		Mat3 c;
#pragma unroll
		for (int i = 0; i < 9; i++) c.m[i] = correct_Rs[9 * (j - 1) + i];
		R = mat3_mul(R, c);
	}
	return R;
}

__global__ void __launch_bounds__(SMPL_THREADS)
smpl_frame_forward_kernel(const __grid_constant__ SmplFrameArgs a)
{
	__shared__ float s_red[SMPL_THREADS / 32][NJ * 3];
	__shared__ float s_pf[NPF + 1];
	__shared__ float s_chain[NJ][12];       // rows 0..2 of the 4x4 chain transform
	__shared__ float s_joint[NJ][3];
	__shared__ float s_rot[NJ][9];
	__shared__ bool s_last;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

	if ((int)blockIdx.x >= a.shape_blocks) {
		// ---- pose blocks: pose_offsets = pose_feature @ posedirs^T, rows [row0, row0 + SMPL_ROWS_PER_BLOCK) ---------------
		if (!a.poseoff) return;
		if (tid < NJ - 1) {
			const Mat3 R = joint_rotation(a.poses, a.correct_Rs, tid + 1);
#pragma unroll
			for (int i = 0; i < 9; i++) s_pf[9 * tid + i] = R.m[i] - ((i % 4 == 0) ? 1.f : 0.f);
		}
		__syncthreads();
		float pf[7];
#pragma unroll
		for (int k = 0; k < 7; k++) pf[k] = (lane + 32 * k < NPF) ? s_pf[lane + 32 * k] : 0.f;
		const int rows = a.V * 3;
		const int row0 = ((int)blockIdx.x - a.shape_blocks) * SMPL_ROWS_PER_BLOCK + warp * (SMPL_ROWS_PER_BLOCK / 8);
#pragma unroll
		for (int r2 = 0; r2 < SMPL_ROWS_PER_BLOCK / 8; r2 += 2) {
			const int ra = row0 + r2, rb = ra + 1;
			float xa[7], xb[7];
#pragma unroll
			for (int k = 0; k < 7; k++) {         // both rows' loads are issued before either is used
				const int c = lane + 32 * k;
				xa[k] = (ra < rows && c < NPF) ? __ldg(a.posedirs + (size_t)ra * NPF + c) : 0.f;
				xb[k] = (rb < rows && c < NPF) ? __ldg(a.posedirs + (size_t)rb * NPF + c) : 0.f;
			}
			float sa = 0.f, sb = 0.f;
#pragma unroll
			for (int k = 0; k < 7; k++) { sa = fmaf(xa[k], pf[k], sa); sb = fmaf(xb[k], pf[k], sb); }
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) { sa += __shfl_xor_sync(0xffffffffu, sa, o); sb += __shfl_xor_sync(0xffffffffu, sb, o); }
			if (lane == 0) {
				if (ra < rows) a.poseoff[ra] = sa;
				if (rb < rows) a.poseoff[rb] = sb;
			}
		}
		return;
	}

	// ---- shape blocks: one thread per vertex ---------------------------------------------------------------------------
	const int v = blockIdx.x * SMPL_THREADS + tid;
	float vs[3] = {0.f, 0.f, 0.f};
	if (v < a.V) {
#pragma unroll
		for (int c = 0; c < 3; c++) {
			float o = 0.f;
			const float* sd = a.shapedirs + ((size_t)v * 3 + c) * a.shape_stride;
			for (int k = 0; k < a.n_shapes; k++) o = fmaf(sd[k], a.shapes[k], o);
			if (a.shapeoff) a.shapeoff[3 * v + c] = o;
			vs[c] = a.v_template[3 * v + c] + o;
		}
	}
	// partial joints of this block: J[j] += J_regressor[j][v] * v_shaped[v]
	for (int j = 0; j < NJ; j++) {
		const float w = v < a.V ? a.J_regressor[(size_t)j * a.V + v] : 0.f;
		float p0 = w * vs[0], p1 = w * vs[1], p2 = w * vs[2];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			p0 += __shfl_xor_sync(0xffffffffu, p0, o); p1 += __shfl_xor_sync(0xffffffffu, p1, o); p2 += __shfl_xor_sync(0xffffffffu, p2, o);
		}
		if (lane == 0) { s_red[warp][3 * j] = p0; s_red[warp][3 * j + 1] = p1; s_red[warp][3 * j + 2] = p2; }
	}
	__syncthreads();
	if (tid < NJ * 3) {
		float s = 0.f;
#pragma unroll
		for (int w = 0; w < SMPL_THREADS / 32; w++) s += s_red[w][tid];
		a.partial[(size_t)blockIdx.x * (NJ * 3) + tid] = s;
	}
	__threadfence();
	__syncthreads();
	if (tid == 0) s_last = atomicAdd(a.ticket, 1u) == (unsigned)a.shape_blocks - 1u;
	__syncthreads();
	if (!s_last) return;
	__threadfence();
	// ---- last shape block: joints (partials in block order), rotations, kinematic chain ----------------------------------
	if (tid < NJ * 3) {
		float s = 0.f;
		for (int b = 0; b < a.shape_blocks; b++) s += __ldcg(a.partial + (size_t)b * (NJ * 3) + tid);
		s_joint[tid / 3][tid % 3] = s;
		if (a.joints) a.joints[tid] = s;
	}
	if (tid >= 96 && tid < 96 + NJ) {           // a different warp than the joint sums: the 24 Rodrigues evaluations
		const int j = tid - 96;
		const Mat3 R = joint_rotation(a.poses, a.correct_Rs, j);
#pragma unroll
		for (int i = 0; i < 9; i++) { s_rot[j][i] = R.m[i]; if (a.rot_mats) a.rot_mats[9 * j + i] = R.m[i]; }
	}
	__syncthreads();
	if (tid == 0) {
		*a.ticket = 0u;
		// the chain is 24 dependent 3x4 products: one thread walks it (parents precede children in SMPL's ordering)
		for (int j = 0; j < NJ; j++) {
			Mat3 R;
#pragma unroll
			for (int i = 0; i < 9; i++) R.m[i] = s_rot[j][i];
			const int p = j == 0 ? -1 : a.parents[j];
			float rel[3];
#pragma unroll
			for (int c = 0; c < 3; c++) rel[c] = s_joint[j][c] - (p >= 0 ? s_joint[p][c] : 0.f);
			if (p < 0) {
#pragma unroll
				for (int r = 0; r < 3; r++) {
					s_chain[j][4 * r] = R.m[3 * r]; s_chain[j][4 * r + 1] = R.m[3 * r + 1]; s_chain[j][4 * r + 2] = R.m[3 * r + 2];
					s_chain[j][4 * r + 3] = rel[r];
				}
			} else {
#pragma unroll
				for (int r = 0; r < 3; r++) {
					const float g0 = s_chain[p][4 * r], g1 = s_chain[p][4 * r + 1], g2 = s_chain[p][4 * r + 2], gt = s_chain[p][4 * r + 3];
#pragma unroll
					for (int c = 0; c < 3; c++) s_chain[j][4 * r + c] = g0 * R.m[c] + g1 * R.m[3 + c] + g2 * R.m[6 + c];
					s_chain[j][4 * r + 3] = g0 * rel[0] + g1 * rel[1] + g2 * rel[2] + gt;
				}
			}
		}
	}
	__syncthreads();
	// A[j] = chain[j] with translation t - R_chain . joint_j (:989-993); bottom row (0,0,0,1)
	if (tid < NJ) {
		const int j = tid;
#pragma unroll
		for (int r = 0; r < 3; r++) {
			const float g0 = s_chain[j][4 * r], g1 = s_chain[j][4 * r + 1], g2 = s_chain[j][4 * r + 2];
			a.A[16 * j + 4 * r] = g0; a.A[16 * j + 4 * r + 1] = g1; a.A[16 * j + 4 * r + 2] = g2;
			a.A[16 * j + 4 * r + 3] = s_chain[j][4 * r + 3] - (g0 * s_joint[j][0] + g1 * s_joint[j][1] + g2 * s_joint[j][2]);
		}
		a.A[16 * j + 12] = 0.f; a.A[16 * j + 13] = 0.f; a.A[16 * j + 14] = 0.f; a.A[16 * j + 15] = 1.f;
	}
}

struct SmplFrameBwdArgs {
	int V;
	const float* posedirs;      // [V*3, 207]
	const int*   parents;       // [24]
	const float* poses;         // [72]
	const float* correct_Rs;    // [23,9] or null
	const float* A;             // [24,16] forward output (chain rotation = A's rotation block)
	const float* joints;        // [24,3]  forward by-product (v_shaped regressed), recomputed by the host-side forward
	const float* d_A;           // [24,16] or null
	const float* d_poseoff;     // [V,3] or null
	float* d_poses;             // [72]
	float* d_correct_Rs;        // [23,9] or null
	float* partial;             // scratch: [pose_blocks][208]
	unsigned* ticket;
	int pose_blocks;
};

__global__ void __launch_bounds__(SMPL_THREADS)
smpl_frame_backward_kernel(const __grid_constant__ SmplFrameBwdArgs a)
{
	__shared__ float s_part[SMPL_THREADS / 32][NPF + 1];
	__shared__ float s_dpf[NPF + 1];
	__shared__ float s_dG[NJ][12];
	__shared__ float s_dR[NJ][9];
	__shared__ bool s_last;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	// ---- d(pose feature)[c] = sum_rows d_poseoff[row] * posedirs[row][c]: this block's rows --------------------------------
	float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
	if (a.d_poseoff) {
		const int rows = a.V * 3;
		const int row0 = (int)blockIdx.x * SMPL_ROWS_PER_BLOCK + warp * (SMPL_ROWS_PER_BLOCK / 8);
#pragma unroll
		for (int r2 = 0; r2 < SMPL_ROWS_PER_BLOCK / 8; r2 += 2) {
			const int ra = row0 + r2, rb = ra + 1;
			const float ga = ra < rows ? a.d_poseoff[ra] : 0.f, gb = rb < rows ? a.d_poseoff[rb] : 0.f;
			float xa[7], xb[7];
#pragma unroll
			for (int k = 0; k < 7; k++) {
				const int c = lane + 32 * k;
				xa[k] = (ra < rows && c < NPF) ? __ldg(a.posedirs + (size_t)ra * NPF + c) : 0.f;
				xb[k] = (rb < rows && c < NPF) ? __ldg(a.posedirs + (size_t)rb * NPF + c) : 0.f;
			}
#pragma unroll
			for (int k = 0; k < 7; k++) acc[k] = fmaf(ga, xa[k], fmaf(gb, xb[k], acc[k]));
		}
	}
#pragma unroll
	for (int k = 0; k < 7; k++)
		if (lane + 32 * k < NPF) s_part[warp][lane + 32 * k] = acc[k];
	__syncthreads();
	if (tid < NPF) {
		float s = 0.f;
#pragma unroll
		for (int w = 0; w < SMPL_THREADS / 32; w++) s += s_part[w][tid];
		a.partial[(size_t)blockIdx.x * (NPF + 1) + tid] = s;
	}
	__threadfence();
	__syncthreads();
	if (tid == 0) s_last = atomicAdd(a.ticket, 1u) == (unsigned)a.pose_blocks - 1u;
	__syncthreads();
	if (!s_last) return;
	__threadfence();
	if (tid < NPF) {
		float s = 0.f;
		for (int b = 0; b < a.pose_blocks; b++) s += __ldcg(a.partial + (size_t)b * (NPF + 1) + tid);      // block order: reproducible
		s_dpf[tid] = s;
	}
	// dL/dG (chain transforms, rows 0..2) from dL/dA: A.R = G.R, A.t = G.t - G.R joint  =>  dG.R = dA.R - dA.t (x) joint, dG.t = dA.t
	if (tid < NJ) {
		const int j = tid;
#pragma unroll
		for (int r = 0; r < 3; r++) {
			const float dt = a.d_A ? a.d_A[16 * j + 4 * r + 3] : 0.f;
#pragma unroll
			for (int c = 0; c < 3; c++) s_dG[j][4 * r + c] = (a.d_A ? a.d_A[16 * j + 4 * r + c] : 0.f) - dt * a.joints[3 * j + c];
			s_dG[j][4 * r + 3] = dt;
		}
	}
	__syncthreads();
	if (tid == 0) {
		*a.ticket = 0u;
		// children before parents: G_j = G_p M_j, M_j = [R_j | rel_j]  =>  dR_j = G_p.R^T dG_j.R,  dG_p.R += dG_j.R R_j^T + dG_j.t (x) rel_j,
		// dG_p.t += dG_j.t.  G_p.R is A_p's rotation block; the root's "parent" is the identity.
		for (int j = NJ - 1; j >= 0; j--) {
			const int p = j == 0 ? -1 : a.parents[j];
			const Mat3 R = joint_rotation(a.poses, a.correct_Rs, j);
			float dGr[9], dGt[3];
#pragma unroll
			for (int r = 0; r < 3; r++) { dGr[3 * r] = s_dG[j][4 * r]; dGr[3 * r + 1] = s_dG[j][4 * r + 1]; dGr[3 * r + 2] = s_dG[j][4 * r + 2]; dGt[r] = s_dG[j][4 * r + 3]; }
			if (p < 0) {
#pragma unroll
				for (int i = 0; i < 9; i++) s_dR[j][i] = dGr[i];
			} else {
				float Gp[9], rel[3];
#pragma unroll
				for (int r = 0; r < 3; r++) { Gp[3 * r] = a.A[16 * p + 4 * r]; Gp[3 * r + 1] = a.A[16 * p + 4 * r + 1]; Gp[3 * r + 2] = a.A[16 * p + 4 * r + 2]; }
#pragma unroll
				for (int c = 0; c < 3; c++) rel[c] = a.joints[3 * j + c] - a.joints[3 * p + c];
#pragma unroll
				for (int r = 0; r < 3; r++)
#pragma unroll
					for (int c = 0; c < 3; c++) {
						s_dR[j][3 * r + c] = Gp[r] * dGr[c] + Gp[3 + r] * dGr[3 + c] + Gp[6 + r] * dGr[6 + c];
						s_dG[p][4 * r + c] += dGr[3 * r] * R.m[3 * c] + dGr[3 * r + 1] * R.m[3 * c + 1] + dGr[3 * r + 2] * R.m[3 * c + 2] + dGt[r] * rel[c];
					}
#pragma unroll
				for (int r = 0; r < 3; r++) s_dG[p][4 * r + 3] += dGt[r];
			}
		}
	}
	__syncthreads();
	// per joint: add the pose-feature gradient, undo the correction product, back-propagate the Rodrigues formula
	if (tid < NJ) {
		const int j = tid;
		Mat3 dR;
#pragma unroll
		for (int i = 0; i < 9; i++) dR.m[i] = s_dR[j][i] + (j >= 1 ? s_dpf[9 * (j - 1) + i] : 0.f);
		const float rx = a.poses[3 * j], ry = a.poses[3 * j + 1], rz = a.poses[3 * j + 2];
		const Mat3 Rod = rodrigues(rx, ry, rz);
		if (a.correct_Rs && j >= 1) {
			Mat3 c;
#pragma unroll
			for (int i = 0; i < 9; i++) c.m[i] = a.correct_Rs[9 * (j - 1) + i];
			if (a.d_correct_Rs) {
				const Mat3 dc = mat3_mul_tn(Rod, dR);          // Rod^T dR
#pragma unroll
				for (int i = 0; i < 9; i++) a.d_correct_Rs[9 * (j - 1) + i] = dc.m[i];
			}
			dR = mat3_mul_nt(dR, c);                            // dR c^T
		} else if (a.d_correct_Rs && j >= 1) {
#pragma unroll
			for (int i = 0; i < 9; i++) a.d_correct_Rs[9 * (j - 1) + i] = 0.f;
		}
		// R = I + s K + (1 - c) K K,  K from k = r / angle,  angle = |r + 1e-8|
		const float ex = rx + 1e-8f, ey = ry + 1e-8f, ez = rz + 1e-8f;
		const float angle = sqrtf(ex * ex + ey * ey + ez * ez);
		const float kx = rx / angle, ky = ry / angle, kz = rz / angle;
		float s, c;
		sincosf(angle, &s, &c);
		const Mat3 K = {{0.f, -kz, ky, kz, 0.f, -kx, -ky, kx, 0.f}};
		const Mat3 KK = mat3_mul(K, K);
		float dangle = 0.f;
#pragma unroll
		for (int i = 0; i < 9; i++) dangle += dR.m[i] * (c * K.m[i] + s * KK.m[i]);
		const Mat3 a1 = mat3_mul_nt(dR, K), a2 = mat3_mul_tn(K, dR);      // d(KK): dR K^T + K^T dR
		Mat3 dK;
#pragma unroll
		for (int i = 0; i < 9; i++) dK.m[i] = s * dR.m[i] + (1.f - c) * (a1.m[i] + a2.m[i]);
		const float dkx = dK.m[7] - dK.m[5], dky = dK.m[2] - dK.m[6], dkz = dK.m[3] - dK.m[1];
		const float inv = 1.f / angle;
		dangle -= (dkx * rx + dky * ry + dkz * rz) * inv * inv;
		a.d_poses[3 * j] = dkx * inv + dangle * ex * inv;
		a.d_poses[3 * j + 1] = dky * inv + dangle * ey * inv;
		a.d_poses[3 * j + 2] = dkz * inv + dangle * ez * inv;
	}
}

} // namespace moss
