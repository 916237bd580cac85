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
// reference.  Here: ONE launch forward, ONE backward, and no dependent global round trip that can be avoided (at this
// size a kernel is a chain of latencies: every fence / ticket / dependent load costs about a microsecond).
//
// Joint regression without a pass over the vertices: joints = J_regressor . (v_template + shapedirs . betas)
//   = J_template + J_shapedirs . betas with J_template = J_regressor . v_template [24,3] and J_shapedirs =
//   J_regressor . shapedirs [24,3,S], two tables of the SMPL model computed ONCE (smpl_joint_tables_kernel).
//
// Forward, grid = 1 chain block + shape blocks + pose blocks, all independent of each other:
//   chain block: joints from the two small tables, 24 Rodrigues evaluations (one thread each), the kinematic chain
//     walked by twelve lanes of one warp out of shared memory -> A [24,4,4], rot_mats, joints;
//   shape blocks (one thread per SMPL vertex): shape_offset rows;
//   pose blocks: 32 consecutive posedirs rows are one contiguous 26.5 KB span; one thread posts it to shared memory as a
//     single bulk asynchronous copy (cp.async.bulk + mbarrier, the 1-D TMA path), meanwhile another warp rebuilds the 207
//     pose features (23 Rodrigues evaluations -- cheaper than a dependent launch); then one warp per row takes the dot
//     products out of shared memory.  All of posedirs (17 MB) is in flight at once: HBM-bound.
// Backward (gradients w.r.t. poses and correct_Rs from dL/dA and dL/dpose_offsets; shapes and the SMPL tables are
// constants of a run), grid = 1 chain block + pose blocks:
//   chain block: back-propagates the chain (children before parents) -> dL/dR of every joint, to scratch;
//   pose blocks: 96 consecutive rows each (three bulk copies posted at once), partial d(pose feature) column sums (the
//     transposed matvec), to scratch;
//   the LAST block to finish (ticket) adds the ~216 partials in a fixed order (reproducible), adds the chain's dL/dR,
//     undoes the correction product and back-propagates the Rodrigues formula.
#pragma once
#include "common.cuh"
#include "skin.cuh"

namespace moss {

constexpr int SMPL_THREADS = 256;
constexpr int SMPL_ROWS_PER_BLOCK = 32;      // posedirs rows per bulk copy (8 warps x 4 rows): 32 * 828 B, a multiple of 16
constexpr int SMPL_BWD_STAGES = 3;           // backward: bulk copies (of SMPL_ROWS_PER_BLOCK rows) per pose block
constexpr int NPF = (NJ - 1) * 9;            // 207 pose features
constexpr int SMPL_MAX_SHAPES = 16;          // betas handled with independent loads (SMPL has 10); more fall back to a loop

// ---- bulk asynchronous copy global -> shared, completion on an mbarrier (PTX ISA: cp.async.bulk, mbarrier) -------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, unsigned bytes, unsigned long long* bar)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		:: "r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
	asm volatile(
		"{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}"
		:: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Rows [row0, row0 + n) of posedirs into s_rows: one bulk copy of the 16-byte-multiple part of the span (thread 0 posts
// it; the barrier must be initialised) plus at most three trailing floats by plain loads (the ragged last span of the
// table: 30 rows of 828 bytes); a table that is not 16-byte aligned is copied by plain loads, eight in flight per thread.
// Returns whether the caller has to wait on the barrier.
__device__ __forceinline__ bool stage_rows(float* s_rows, unsigned long long* bar, const float* posedirs, int row0, int n)
{
	const float* src = posedirs + (size_t)row0 * NPF;
	const int total = n * NPF;
	if (!((size_t)src & 15)) {
		const int body = total & ~3;
		if (threadIdx.x == 0) bulk_load(s_rows, src, (unsigned)body * 4u, bar);
		if ((int)threadIdx.x < total - body) s_rows[body + threadIdx.x] = __ldg(src + body + threadIdx.x);
		return true;
	}
	for (int i0 = threadIdx.x; i0 < total; i0 += 8 * blockDim.x) {
		float x[8];
#pragma unroll
		for (int k = 0; k < 8; k++) x[k] = i0 + k * blockDim.x < total ? __ldg(src + i0 + k * blockDim.x) : 0.f;
#pragma unroll
		for (int k = 0; k < 8; k++) if (i0 + k * blockDim.x < total) s_rows[i0 + k * blockDim.x] = x[k];
	}
	return false;
}

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

// rotation of joint j after the optional correction (:874-877): thread-local, used by every block role
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

// ---- one-time tables of the SMPL model: J_template = J_regressor . v_template, J_shapedirs = J_regressor . shapedirs -----
// grid (24, 1 + S): block (j, 0) -> J_template[j][0..2], block (j, 1 + k) -> J_shapedirs[j][0..2][k]; fixed-order tree.
__global__ void __launch_bounds__(SMPL_THREADS)
smpl_joint_tables_kernel(int V, int S, const float* __restrict__ v_template, const float* __restrict__ shapedirs,
	const float* __restrict__ J_regressor, float* __restrict__ J_template, float* __restrict__ J_shapedirs)
{
	__shared__ float s_red[SMPL_THREADS / 32][3];
	const int j = blockIdx.x, k = (int)blockIdx.y - 1;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	float p[3] = {0.f, 0.f, 0.f};
	for (int v = tid; v < V; v += SMPL_THREADS) {
		const float w = J_regressor[(size_t)j * V + v];
#pragma unroll
		for (int c = 0; c < 3; c++)
			p[c] = fmaf(w, k < 0 ? v_template[3 * v + c] : shapedirs[((size_t)v * 3 + c) * S + k], p[c]);
	}
#pragma unroll
	for (int c = 0; c < 3; c++) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) p[c] += __shfl_xor_sync(0xffffffffu, p[c], o);
		if (lane == 0) s_red[warp][c] = p[c];
	}
	__syncthreads();
	if (tid < 3) {
		float s = 0.f;
#pragma unroll
		for (int w = 0; w < SMPL_THREADS / 32; w++) s += s_red[w][tid];
		if (k < 0) J_template[3 * j + tid] = s;
		else J_shapedirs[(size_t)(3 * j + tid) * S + k] = s;
	}
}

struct SmplFrameArgs {
	int V;                      // SMPL vertices (6890)
	int n_shapes;               // betas used (<= shape_stride)
	int shape_stride;           // 10
	const float* shapedirs;     // [V,3,shape_stride]
	const float* posedirs;      // [V*3, 207] (reference layout [V,3,207])
	const float* J_template;    // [24,3]
	const float* J_shapedirs;   // [24,3,shape_stride]
	const int*   parents;       // [24]  (kintree_table[0]; parents[0] ignored)
	const float* poses;         // [72] axis-angle
	const float* shapes;        // [n_shapes]
	const float* correct_Rs;    // [23,9] or null
	float* A;                   // [24,16]
	float* rot_mats;            // [24,9]  (after the pose correction) or null
	float* joints;              // [24,3]  (the backward reads it) or null
	float* poseoff;             // [V,3] or null (no pose blocks then)
	float* shapeoff;            // [V,3] or null (no shape blocks then)
	int shape_blocks;
};

__global__ void __launch_bounds__(SMPL_THREADS)
smpl_frame_forward_kernel(const __grid_constant__ SmplFrameArgs a)
{
	__shared__ __align__(128) float s_rows[SMPL_ROWS_PER_BLOCK * NPF];     // pose blocks: the staged posedirs rows
	__shared__ unsigned long long s_bar;
	__shared__ float s_pf[NPF + 1];
	__shared__ float s_chain[NJ][12];       // rows 0..2 of the 4x4 chain transform
	__shared__ float s_joint[NJ][3];
	__shared__ float s_rot[NJ][9];
	__shared__ float s_rel[NJ][3];
	__shared__ int s_par[NJ];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

	if ((int)blockIdx.x > a.shape_blocks) {
		// ---- pose blocks: pose_offsets = pose_feature @ posedirs^T, rows [row0, row0 + SMPL_ROWS_PER_BLOCK) ---------------
		const int rows = a.V * 3;
		const int row0 = ((int)blockIdx.x - a.shape_blocks - 1) * SMPL_ROWS_PER_BLOCK;
		const int n = min(SMPL_ROWS_PER_BLOCK, rows - row0);
		if (tid == 0) { mbar_init(&s_bar, 1); mbar_init_fence(); }
		const bool bulk = stage_rows(s_rows, &s_bar, a.posedirs, row0, n);
		if (tid >= 32 && tid < 32 + NJ - 1) {      // not the warp that posted the copy
			const int j = tid - 32;
			const Mat3 R = joint_rotation(a.poses, a.correct_Rs, j + 1);
#pragma unroll
			for (int i = 0; i < 9; i++) s_pf[9 * j + i] = R.m[i] - ((i % 4 == 0) ? 1.f : 0.f);
		}
		__syncthreads();
		if (bulk) mbar_wait(&s_bar, 0);
		float pf[7];
#pragma unroll
		for (int k = 0; k < 7; k++) pf[k] = (lane + 32 * k < NPF) ? s_pf[lane + 32 * k] : 0.f;
#pragma unroll
		for (int r = 0; r < SMPL_ROWS_PER_BLOCK / 8; r++) {
			const int lr = warp * (SMPL_ROWS_PER_BLOCK / 8) + r;
			float sa = 0.f;
			if (lr < n) {
#pragma unroll
				for (int k = 0; k < 7; k++) {
					const int c = lane + 32 * k;
					if (c < NPF) sa = fmaf(s_rows[lr * NPF + c], pf[k], sa);
				}
			}
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) sa += __shfl_xor_sync(0xffffffffu, sa, o);
			if (lane == 0 && lr < n) a.poseoff[row0 + lr] = sa;
		}
		return;
	}

	if (blockIdx.x >= 1) {
		// ---- shape blocks: shape_offset[v][c] = shapedirs[v][c][:] . betas, one thread per vertex --------------------------
		const int v = ((int)blockIdx.x - 1) * SMPL_THREADS + tid;
		if (v >= a.V) return;
#pragma unroll
		for (int c = 0; c < 3; c++) {
			const float* sd = a.shapedirs + ((size_t)v * 3 + c) * a.shape_stride;
			float x[SMPL_MAX_SHAPES];
#pragma unroll
			for (int k = 0; k < SMPL_MAX_SHAPES; k++) x[k] = k < a.n_shapes ? __ldg(sd + k) : 0.f;   // independent loads, one round trip
			float o = 0.f;
#pragma unroll
			for (int k = 0; k < SMPL_MAX_SHAPES; k++) if (k < a.n_shapes) o = fmaf(x[k], a.shapes[k], o);
			for (int k = SMPL_MAX_SHAPES; k < a.n_shapes; k++) o = fmaf(sd[k], a.shapes[k], o);
			a.shapeoff[3 * v + c] = o;
		}
		return;
	}

	// ---- chain block: joints, rotations, kinematic chain -----------------------------------------------------------------
	if (tid < NJ * 3) {             // joints[j][c] = J_template + J_shapedirs . betas
		const float* js = a.J_shapedirs + (size_t)tid * a.shape_stride;
		float x[SMPL_MAX_SHAPES];
#pragma unroll
		for (int k = 0; k < SMPL_MAX_SHAPES; k++) x[k] = k < a.n_shapes ? __ldg(js + k) : 0.f;
		float s = 0.f;
#pragma unroll
		for (int k = 0; k < SMPL_MAX_SHAPES; k++) if (k < a.n_shapes) s = fmaf(x[k], a.shapes[k], s);
		for (int k = SMPL_MAX_SHAPES; k < a.n_shapes; k++) s = fmaf(js[k], a.shapes[k], s);
		s += a.J_template[tid];
		s_joint[tid / 3][tid % 3] = s;
		if (a.joints) a.joints[tid] = s;
	}
	if (tid >= 96 && tid < 96 + NJ) {           // a different warp than the joints: the 24 Rodrigues evaluations
		const int j = tid - 96;
		s_par[j] = j == 0 ? -1 : a.parents[j];  // no global load may sit on the serial walk below
		const Mat3 R = joint_rotation(a.poses, a.correct_Rs, j);
#pragma unroll
		for (int i = 0; i < 9; i++) { s_rot[j][i] = R.m[i]; if (a.rot_mats) a.rot_mats[9 * j + i] = R.m[i]; }
	}
	__syncthreads();
	if (tid < NJ * 3) {             // bone vectors: rel_j = joint_j - joint_parent(j) (:973-974)
		const int j = tid / 3, c = tid % 3, p = s_par[j];
		s_rel[j][c] = s_joint[j][c] - (p >= 0 ? s_joint[p][c] : 0.f);
	}
	__syncthreads();
	if (warp == 0) {
		// The chain is 24 dependent 3x4 products G_j = G_parent . [R_j | rel_j] (parents precede children in SMPL's ordering,
		// :983-987): twelve lanes hold one element each, one shared-memory round trip + four FMAs per joint.
		const int r = lane >> 2, c = lane & 3;
		for (int j = 0; j < NJ; j++) {
			const int p = s_par[j];
			if (lane < 12) {
				const float m0 = c < 3 ? s_rot[j][c] : s_rel[j][0], m1 = c < 3 ? s_rot[j][3 + c] : s_rel[j][1],
					m2 = c < 3 ? s_rot[j][6 + c] : s_rel[j][2];
				float v;
				if (p < 0) v = r == 0 ? m0 : (r == 1 ? m1 : m2);
				else {
					const float g0 = s_chain[p][4 * r], g1 = s_chain[p][4 * r + 1], g2 = s_chain[p][4 * r + 2];
					v = g0 * m0 + g1 * m1 + g2 * m2;
					if (c == 3) v += s_chain[p][4 * r + 3];
				}
				s_chain[j][4 * r + c] = v;
			}
			__syncwarp();
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
	const float* joints;        // [24,3]  forward output
	const float* d_A;           // [24,16] or null
	const float* d_poseoff;     // [V,3] or null (no pose blocks then)
	float* d_poses;             // [72]
	float* d_correct_Rs;        // [23,9] or null
	float* partial;             // scratch: [pose_blocks][208]
	float* chain_dR;            // scratch: [24][9]
	unsigned* ticket;           // scratch: zero on entry, zero on exit
	int pose_blocks;
};

__global__ void __launch_bounds__(SMPL_THREADS)
smpl_frame_backward_kernel(const __grid_constant__ SmplFrameBwdArgs a)
{
	extern __shared__ __align__(128) float s_dyn[];     // pose blocks: SMPL_BWD_STAGES spans of SMPL_ROWS_PER_BLOCK rows
	__shared__ unsigned long long s_bar[SMPL_BWD_STAGES];
	__shared__ float s_g[SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK];
	__shared__ float s_part[SMPL_THREADS / 32][NPF + 1];
	__shared__ float s_dG[NJ][12];
	__shared__ float s_dR[NJ][9];
	__shared__ float s_rot[NJ][9];
	__shared__ float s_Ar[NJ][9];
	__shared__ float s_jt[NJ][3];
	__shared__ int s_par[NJ];
	__shared__ bool s_last;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

	if (blockIdx.x >= 1) {
		// ---- pose blocks: d(pose feature)[c] = sum_rows d_poseoff[row] * posedirs[row][c], one column per thread -----------
		const int rows = a.V * 3;
		const int row0 = ((int)blockIdx.x - 1) * (SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK);
		if (tid == 0) {
#pragma unroll
			for (int s = 0; s < SMPL_BWD_STAGES; s++) mbar_init(&s_bar[s], 1);
			mbar_init_fence();
		}
		bool bulk[SMPL_BWD_STAGES];
		int n[SMPL_BWD_STAGES];
#pragma unroll
		for (int s = 0; s < SMPL_BWD_STAGES; s++) {        // all spans are posted before anything waits
			n[s] = max(0, min(SMPL_ROWS_PER_BLOCK, rows - (row0 + s * SMPL_ROWS_PER_BLOCK)));
			bulk[s] = n[s] > 0 && stage_rows(s_dyn + s * SMPL_ROWS_PER_BLOCK * NPF, &s_bar[s], a.posedirs, row0 + s * SMPL_ROWS_PER_BLOCK, n[s]);
		}
		if (tid >= 32 && tid < 32 + SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK)
			s_g[tid - 32] = (row0 + tid - 32 < rows) ? a.d_poseoff[row0 + tid - 32] : 0.f;
		__syncthreads();
		float acc = 0.f, acc2 = 0.f;
#pragma unroll
		for (int s = 0; s < SMPL_BWD_STAGES; s++) {
			if (bulk[s]) mbar_wait(&s_bar[s], 0);
			if (tid < NPF) {
				const float* sr = s_dyn + s * SMPL_ROWS_PER_BLOCK * NPF + tid;
				const float* g = s_g + s * SMPL_ROWS_PER_BLOCK;
#pragma unroll 8
				for (int r = 0; r + 1 < n[s]; r += 2) {
					acc = fmaf(g[r], sr[r * NPF], acc);
					acc2 = fmaf(g[r + 1], sr[(r + 1) * NPF], acc2);
				}
				if (n[s] & 1) acc = fmaf(g[n[s] - 1], sr[(n[s] - 1) * NPF], acc);
			}
		}
		if (tid < NPF) a.partial[(size_t)(blockIdx.x - 1) * (NPF + 1) + tid] = acc + acc2;
	} else {
		// ---- chain block: dL/dR of every joint from dL/dA --------------------------------------------------------------------
		// everything the serial walk reads goes to shared memory first: no global load may sit on that path
		if (tid < NJ) {
			const int j = tid;
			const Mat3 R = joint_rotation(a.poses, a.correct_Rs, j);
#pragma unroll
			for (int i = 0; i < 9; i++) s_rot[j][i] = R.m[i];
		} else if (tid >= 32 && tid < 32 + NJ) {
			const int j = tid - 32;
			s_par[j] = j == 0 ? -1 : a.parents[j];
			float jt[3];
#pragma unroll
			for (int c = 0; c < 3; c++) { jt[c] = a.joints[3 * j + c]; s_jt[j][c] = jt[c]; }
			// dL/dG (chain transforms, rows 0..2) from dL/dA: A.R = G.R, A.t = G.t - G.R joint  =>  dG.R = dA.R - dA.t (x) joint, dG.t = dA.t
#pragma unroll
			for (int r = 0; r < 3; r++) {
#pragma unroll
				for (int c = 0; c < 3; c++) s_Ar[j][3 * r + c] = a.A[16 * j + 4 * r + c];
				const float dt = a.d_A ? a.d_A[16 * j + 4 * r + 3] : 0.f;
#pragma unroll
				for (int c = 0; c < 3; c++) s_dG[j][4 * r + c] = (a.d_A ? a.d_A[16 * j + 4 * r + c] : 0.f) - dt * jt[c];
				s_dG[j][4 * r + 3] = dt;
			}
		}
		__syncthreads();
		if (warp == 0) {
			// children before parents: G_j = G_p M_j, M_j = [R_j | rel_j]  =>  dR_j = G_p.R^T dG_j.R,  dG_p.R += dG_j.R R_j^T + dG_j.t (x) rel_j,
			// dG_p.t += dG_j.t.  G_p.R is A_p's rotation block; the root's "parent" is the identity.  Lanes 0..8 own one
			// element (r,c) of the two 3x3 results, lanes 9..11 one translation component.
			const int r = lane / 3, c = lane % 3;
			for (int j = NJ - 1; j >= 0; j--) {
				const int p = s_par[j];
				if (lane < 9) {
					if (p < 0) s_dR[j][3 * r + c] = s_dG[j][4 * r + c];
					else {
						s_dR[j][3 * r + c] = s_Ar[p][r] * s_dG[j][c] + s_Ar[p][3 + r] * s_dG[j][4 + c] + s_Ar[p][6 + r] * s_dG[j][8 + c];
						s_dG[p][4 * r + c] += s_dG[j][4 * r] * s_rot[j][3 * c] + s_dG[j][4 * r + 1] * s_rot[j][3 * c + 1] +
							s_dG[j][4 * r + 2] * s_rot[j][3 * c + 2] + s_dG[j][4 * r + 3] * (s_jt[j][c] - s_jt[p][c]);
					}
				} else if (lane < 12 && p >= 0) {
					s_dG[p][4 * (lane - 9) + 3] += s_dG[j][4 * (lane - 9) + 3];
				}
				__syncwarp();
			}
		}
		__syncthreads();
		if (tid < NJ * 9) a.chain_dR[tid] = s_dR[tid / 9][tid % 9];
	}

	// ---- the last block to finish: partials in a fixed order, the chain's dL/dR, correction product, Rodrigues backward ---
	__threadfence();
	__syncthreads();
	if (tid == 0) s_last = atomicAdd(a.ticket, 1u) == (unsigned)a.pose_blocks;        // pose_blocks + the chain block
	__syncthreads();
	if (!s_last) return;
	__threadfence();
	{
		// warp w adds blocks w, w+8, ... (7 columns per lane, all loads independent), then the eight warp sums are added in order
		float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
		for (int b0 = warp; b0 < a.pose_blocks; b0 += 4 * (SMPL_THREADS / 32)) {
			float x[4][7];
#pragma unroll
			for (int i = 0; i < 4; i++)
#pragma unroll
				for (int k = 0; k < 7; k++) {
					const int b = b0 + i * (SMPL_THREADS / 32), c = lane + 32 * k;
					x[i][k] = (b < a.pose_blocks && c < NPF) ? __ldcg(a.partial + (size_t)b * (NPF + 1) + c) : 0.f;
				}
#pragma unroll
			for (int i = 0; i < 4; i++)
#pragma unroll
				for (int k = 0; k < 7; k++) acc[k] += x[i][k];
		}
#pragma unroll
		for (int k = 0; k < 7; k++)
			if (lane + 32 * k < NPF) s_part[warp][lane + 32 * k] = acc[k];
	}
	__syncthreads();
	if (tid == 0) *a.ticket = 0u;
	// per joint: chain + pose-feature gradient, undo the correction product, back-propagate the Rodrigues formula
	if (tid < NJ) {
		const int j = tid;
		Mat3 dR;
#pragma unroll
		for (int i = 0; i < 9; i++) {
			float s = 0.f;
			if (j >= 1) {
#pragma unroll
				for (int w = 0; w < SMPL_THREADS / 32; w++) s += s_part[w][9 * (j - 1) + i];
			}
			dR.m[i] = __ldcg(a.chain_dR + 9 * j + i) + s;
		}
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

static_assert((SMPL_ROWS_PER_BLOCK * NPF * 4) % 16 == 0, "a span of rows is a 16-byte multiple");
static_assert(32 + SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK <= SMPL_THREADS, "one thread per staged gradient row");
static_assert(NJ * 9 <= SMPL_THREADS, "one thread per chain gradient element");
constexpr size_t SMPL_BWD_DYN_SMEM = (size_t)SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK * NPF * sizeof(float);

} // namespace moss
