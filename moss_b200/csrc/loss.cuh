// Image-space losses fused behind the rasterizer.  What train_ZJU.py:107-128 computes
// from the rendered image and alpha with ~60 torch launches, three boolean-mask gathers (each a host sync for the
// output size), five grouped 11x11 convolutions per SSIM call and a ten times wider image for S3IM --
//     Ll1       = l1_loss(image.permute(1,2,0)[bound_mask[0]==1], gt_image.permute(1,2,0)[bound_mask[0]==1])   train_ZJU.py:111
//     mask_loss = l2_loss(alpha[bound_mask==1], bkgd_mask[bound_mask==1])                                      :112
//     ssim_loss = ssim(image[:, y:y+h, x:x+w], gt_image[:, y:y+h, x:x+w])         utils/loss_utils.py:57-87    :115-119
//     s3im_loss = s3im_fun(img_pred, img_gt)                                      utils/loss_utils.py:16-38    :123
//     loss      = Ll1 + 0.5 mask_loss + 0.2 (1 - ssim_loss) + 0.3 s3im_loss  (+ LPIPS and the pose NLL, not here) :131
// -- becomes ONE launch that writes the four loss values and dL/dimage, dL/dalpha (what the rasterizer's backward
// consumes) without materialising a gather, a convolution output or the stretched image.
//
// S3IM with a batch of one is deterministic: randperm(1) is always [0] (loss_utils.py:26-30), so the ten "shuffled"
// copies are the same pixel repeated ten times along x (the permute/reshape at :35-36 interleaves them), and the loss is
// SSIM on that stretched image.  An 11-tap window over a x10 stretch touches at most three source pixels; per replica r
// the horizontal filter collapses to three taps (a_r, b_r, c_r) on (x-1, x, x+1), the vertical one is unchanged.  So
// S3IM = the mean over r of ten SSIM maps that share one vertical pass.
//
// One block per 32x32 pixel tile, per colour channel in shared memory:
//   A  tile + 10-pixel halo of image and ground truth (zero outside the crop: conv2d pads the CROPPED image with zeros)
//   B  horizontal 11-tap pass of (x, y, xx, yy, xy)                 C  vertical pass -> SSIM value and the three partial
//   derivatives of the map (w.r.t. the window means of x, xx, xy) on the tile + 5-pixel halo
//   D/E the two passes again over the derivative maps -> dSSIM/dimage on the tile
//   S  the S3IM variant of B..E (vertical pass first, ten 3-tap horizontal variants)
// then the masked L1 / L2 terms, one block-level sum per loss, and the last block to finish (ticket) adds the per-block
// sums in block order (reproducible) and writes the loss values.
#pragma once
#include "common.cuh"

namespace moss {

constexpr int LOSS_T = 32;                       // output tile edge
constexpr int LOSS_R = 5;                        // window radius (window_size = 11, loss_utils.py:57)
constexpr int LOSS_WIN = 2 * LOSS_R + 1;
constexpr int LOSS_IN = LOSS_T + 4 * LOSS_R;     // 52: input tile edge, two window passes deep
constexpr int LOSS_MID = LOSS_T + 2 * LOSS_R;    // 42: where the statistics and the derivative maps live
// threads per block: a template parameter.  The passes are latency-bound, so with few tiles (ZJU) 16 warps per block hide
// more of it (116 -> 74 us per tile block); with many tiles two resident 8-warp blocks per SM do better (1080p: 328 vs 381 us).
constexpr int LOSS_S3_REP = 10;                  // repeat_time (loss_utils.py:16)
constexpr int LOSS_S3_W = LOSS_T + 2;            // 34: S3IM map points per row (tile + 1 pixel each side)
constexpr int LOSS_S3_VW = LOSS_T + 4;           // 36: columns of the vertical pass the S3IM maps read
// shared memory (floats): the two input tiles, the pass buffer (5 maps of 52x42), the derivative maps (3 of 42x42)
constexpr int LOSS_SM_IN = LOSS_IN * LOSS_IN;
constexpr int LOSS_SM_H = 5 * LOSS_IN * LOSS_MID;
constexpr int LOSS_SM_D = 3 * LOSS_MID * LOSS_MID;
constexpr size_t LOSS_DYN_SMEM = (size_t)(2 * LOSS_SM_IN + LOSS_SM_H + LOSS_SM_D) * sizeof(float);
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;       // loss_utils.py:78-79

struct LossArgs {
	int W, H;
	const float* image;       // [3,H,W]
	const float* alpha;       // [H,W] or null
	const float* gt;          // [3,H,W]
	const float* bkgd;        // [H,W] or null
	const float* bound;       // [H,W] or null (null: every pixel counts)
	const int*   rect;        // device int[5]: crop x, y, w, h and the number of bound == 1 pixels
	float w_l1, w_mask, w_ssim, w_s3im;
	float g[LOSS_WIN];        // the normalised 1-D window (loss_utils.py:47-49)
	float s3[LOSS_S3_REP][3]; // S3IM: taps on (x-1, x, x+1) of replica r
	float* dL_dimage;         // [3,H,W] or null
	float* dL_dalpha;         // [H,W] or null
	float* losses;            // [5]: Ll1, mask_loss, ssim, s3im_loss, weighted total
	float* partial;           // scratch [blocks][4]
	unsigned* ticket;         // scratch, zero on entry and on exit
};

// SSIM map value and its partial derivatives w.r.t. the window means of x (mu1), x*x (e11) and x*y (e12), the means of
// y being constants (loss_utils.py:65-81).  dmu1 is the TOTAL derivative: sigma1_sq = e11 - mu1^2 and sigma12 = e12 -
// mu1 mu2 depend on mu1 as well.
__device__ __forceinline__ float ssim_point(float mu1, float mu2, float e11, float e22, float e12, float& dmu1, float& de11, float& de12)
{
	const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
	const float s11 = e11 - mu1_sq, s22 = e22 - mu2_sq, s12 = e12 - mu12;
	const float A1 = 2.f * mu12 + SSIM_C1, A2 = 2.f * s12 + SSIM_C2;
	const float B1 = mu1_sq + mu2_sq + SSIM_C1, B2 = s11 + s22 + SSIM_C2;
	// two reciprocals (MUFU.RCP + one Newton step each: full fp32 accuracy, a third of the instructions of four divisions)
	float rB1 = __frcp_rn(B1), rB2 = __frcp_rn(B2);
	const float inv = rB1 * rB2;
	const float map = A1 * A2 * inv;
	de12 = 2.f * A1 * inv;
	de11 = -map * rB2;
	dmu1 = 2.f * mu2 * (A2 - A1) * inv + 2.f * mu1 * map * (rB2 - rB1);
	return map;
}

// CSPLIT: one block per (tile, colour channel) instead of one per tile (grid.z = 3).  The three channels of a tile are
// independent up to the sums, so with few tiles (ZJU: 96 tiles inside the crop, fewer than SMs) the split puts three
// times as many blocks on the machine and every block runs a third of the barrier-separated passes; block z = 0 also
// owns the alpha / mask term.  Partial sums stay per block and are added in block order: reproducible.
// Measured on B200 (kernel us, old -> new): 256 x 256 71 -> 43, 512 x 512 (ZJU crop) 75 -> 65.  The gain is bounded by
// quantisation: 288 equal blocks on 148 SMs x 2 resident blocks is one wave whose SMs run two blocks at ~60 % issue
// rate; 64 registers (two resident 16-warp blocks) beat 100 registers (one) at ZJU size, 68 vs 73 us.
// The launch bounds matter for the per-tile kernel too: (256, 1) let ptxas take 172 registers -> ONE resident block
// and 1080p went from 327 to 505 us; (256, 2) gives 116 registers, no spills, two resident blocks.
constexpr int LOSS_CSPLIT_MAX_TILES = 2 * 148;   // split when the launch has at most this many tiles (host side)
#ifndef MOSS_LOSS_SPLIT_THREADS
#define MOSS_LOSS_SPLIT_THREADS 512              // threads of a per-channel block
#endif
#ifndef MOSS_LOSS_SPLIT_MINBLOCKS
#define MOSS_LOSS_SPLIT_MINBLOCKS 2              // resident per-channel blocks per SM the register budget allows
#endif
#ifndef MOSS_LOSS_NOSPLIT_MINBLOCKS
#define MOSS_LOSS_NOSPLIT_MINBLOCKS 2            // per-tile blocks: two resident 8-warp blocks per SM
#endif
constexpr int LOSS_SPLIT_THREADS = MOSS_LOSS_SPLIT_THREADS;

template <int LOSS_THREADS, bool CSPLIT>
__global__ void __launch_bounds__(LOSS_THREADS, CSPLIT ? MOSS_LOSS_SPLIT_MINBLOCKS : MOSS_LOSS_NOSPLIT_MINBLOCKS)
image_losses_kernel(const __grid_constant__ LossArgs a)
{
	constexpr int NCH = CSPLIT ? 1 : 3;                                                       // channels of this block
	const int ch_first = CSPLIT ? (int)blockIdx.z : 0;
	constexpr int LOSS_PIX = LOSS_T * LOSS_T / LOSS_THREADS;                                  // output pixels per thread
	constexpr int LOSS_S3_PTS = (LOSS_MID * LOSS_S3_W + LOSS_THREADS - 1) / LOSS_THREADS;     // S3IM map points per thread
	static_assert(LOSS_T * LOSS_T % LOSS_THREADS == 0 && (LOSS_THREADS / 4) % 32 == 0, "whole pixels per thread, whole warps per loss");
	extern __shared__ __align__(16) float sm[];
	float* s_in1 = sm;                                  // [52][52] image tile, zero outside the crop
	float* s_in2 = s_in1 + LOSS_SM_IN;                  // [52][52] ground truth tile
	float* s_h = s_in2 + LOSS_SM_IN;                    // pass buffer
	float* s_d = s_h + LOSS_SM_H;                       // derivative maps
	__shared__ float s_red[LOSS_THREADS / 32][4];
	__shared__ bool s_last;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int W = a.W, H = a.H;
	const size_t N = (size_t)W * H;
	// the crop, clipped to the image (a rectangle from moss_mask_bounds always lies inside)
	const int rx = max(a.rect[0], 0), ry = max(a.rect[1], 0);
	const int rw = max(min(a.rect[0] + a.rect[2], W) - rx, 0), rh = max(min(a.rect[1] + a.rect[3], H) - ry, 0), count = a.rect[4];
	// The tile grid is aligned to the CROP's corner, not to the image's (the launch has one extra tile column and row for
	// that): a 254 x 383 crop then meets 8 x 12 tiles instead of 8 x 13 with two partial rows -- at ZJU size the difference
	// between one resident wave of per-channel blocks and a wave plus a tail.
	const int tx0 = (int)blockIdx.x * LOSS_T - ((LOSS_T - (rw > 0 && rh > 0 ? rx : 0) % LOSS_T) % LOSS_T);
	const int ty0 = (int)blockIdx.y * LOSS_T - ((LOSS_T - (rw > 0 && rh > 0 ? ry : 0) % LOSS_T) % LOSS_T);
	const bool ssim_on = (a.w_ssim != 0.f || a.w_s3im != 0.f) && rw > 0 && rh > 0 &&
		tx0 < rx + rw && tx0 + LOSS_T > rx && ty0 < ry + rh && ty0 + LOSS_T > ry;
	auto in_rect = [&](int x, int y) { return x >= rx && x < rx + rw && y >= ry && y < ry + rh; };

	float grad[LOSS_PIX][NCH];      // dL/dimage of this thread's pixels: pixel k is tile-local (tid + k*LOSS_THREADS)
#pragma unroll
	for (int k = 0; k < LOSS_PIX; k++)
#pragma unroll
		for (int c = 0; c < NCH; c++) grad[k][c] = 0.f;
	float sum_l1 = 0.f, sum_l2 = 0.f, sum_ssim = 0.f, sum_s3 = 0.f;

	if (ssim_on) {
		const float k_ssim = -a.w_ssim / (3.f * (float)rw * (float)rh);                       // d/dmap of w (1 - mean(map))
		const float k_s3 = -a.w_s3im / (3.f * (float)rw * (float)rh * (float)LOSS_S3_REP);
		for (int c = 0; c < NCH; c++) {
			const int ch = ch_first + c;
			const float* img = a.image + ch * N;
			const float* gt = a.gt + ch * N;
			// ---- A: tile + halo, zero outside the crop -------------------------------------------------------------------
			for (int i = tid; i < LOSS_SM_IN; i += LOSS_THREADS) {
				const int y = ty0 - 2 * LOSS_R + i / LOSS_IN, x = tx0 - 2 * LOSS_R + i % LOSS_IN;
				const bool in = in_rect(x, y);
				s_in1[i] = in ? __ldg(img + (size_t)y * W + x) : 0.f;
				s_in2[i] = in ? __ldg(gt + (size_t)y * W + x) : 0.f;
			}
			__syncthreads();
			if (a.w_ssim != 0.f) {
				// ---- B: horizontal pass of (x, y, xx, yy, xy): rows 0..51, columns 0..41 (centre = input column + 5) ------
				for (int i = tid; i < LOSS_IN * LOSS_MID; i += LOSS_THREADS) {
					const int yy = i / LOSS_MID, xx = i % LOSS_MID;
					float h1 = 0.f, h2 = 0.f, h11 = 0.f, h22 = 0.f, h12 = 0.f;
#pragma unroll
					for (int k = 0; k < LOSS_WIN; k++) {
						const float u = s_in1[yy * LOSS_IN + xx + k], v = s_in2[yy * LOSS_IN + xx + k], w = a.g[k];
						const float wu = w * u, wv = w * v;
						h1 += wu; h2 += wv; h11 = fmaf(wu, u, h11); h22 = fmaf(wv, v, h22); h12 = fmaf(wu, v, h12);
					}
					s_h[i] = h1; s_h[LOSS_IN * LOSS_MID + i] = h2; s_h[2 * LOSS_IN * LOSS_MID + i] = h11;
					s_h[3 * LOSS_IN * LOSS_MID + i] = h22; s_h[4 * LOSS_IN * LOSS_MID + i] = h12;
				}
				__syncthreads();
				// ---- C: vertical pass -> statistics on the tile + 5-pixel halo, map value, derivative maps -----------------
				for (int i = tid; i < LOSS_MID * LOSS_MID; i += LOSS_THREADS) {
					const int yy = i / LOSS_MID, xx = i % LOSS_MID;
					const int qx = tx0 - LOSS_R + xx, qy = ty0 - LOSS_R + yy;
					float dA = 0.f, dB = 0.f, dC = 0.f;
					if (in_rect(qx, qy)) {
						float m[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
						for (int k = 0; k < LOSS_WIN; k++) {
							const float w = a.g[k];
#pragma unroll
							for (int q = 0; q < 5; q++) m[q] = fmaf(w, s_h[q * LOSS_IN * LOSS_MID + (yy + k) * LOSS_MID + xx], m[q]);
						}
						const float map = ssim_point(m[0], m[1], m[2], m[3], m[4], dA, dB, dC);
						if (xx >= LOSS_R && xx < LOSS_R + LOSS_T && yy >= LOSS_R && yy < LOSS_R + LOSS_T) sum_ssim += map;
					}
					s_d[i] = dA; s_d[LOSS_MID * LOSS_MID + i] = dB; s_d[2 * LOSS_MID * LOSS_MID + i] = dC;
				}
				__syncthreads();
				// ---- D: horizontal pass over the derivative maps: rows 0..41, output columns 0..31 --------------------------
				for (int i = tid; i < LOSS_MID * LOSS_T; i += LOSS_THREADS) {
					const int yy = i / LOSS_T, xx = i % LOSS_T;
					float o[3] = {0.f, 0.f, 0.f};
#pragma unroll
					for (int k = 0; k < LOSS_WIN; k++) {
						const float w = a.g[k];
#pragma unroll
						for (int q = 0; q < 3; q++) o[q] = fmaf(w, s_d[q * LOSS_MID * LOSS_MID + yy * LOSS_MID + xx + k], o[q]);
					}
#pragma unroll
					for (int q = 0; q < 3; q++) s_h[q * LOSS_MID * LOSS_T + i] = o[q];
				}
				__syncthreads();
				// ---- E: vertical pass -> dSSIM/dimage of this thread's pixels -------------------------------------------------
#pragma unroll
				for (int k = 0; k < LOSS_PIX; k++) {
					const int p = tid + k * LOSS_THREADS, py = p / LOSS_T, px = p % LOSS_T;
					if (!in_rect(tx0 + px, ty0 + py)) continue;
					float o[3] = {0.f, 0.f, 0.f};
#pragma unroll
					for (int t = 0; t < LOSS_WIN; t++) {
						const float w = a.g[t];
#pragma unroll
						for (int q = 0; q < 3; q++) o[q] = fmaf(w, s_h[q * LOSS_MID * LOSS_T + (py + t) * LOSS_T + px], o[q]);
					}
					const float u = s_in1[(py + 2 * LOSS_R) * LOSS_IN + px + 2 * LOSS_R], v = s_in2[(py + 2 * LOSS_R) * LOSS_IN + px + 2 * LOSS_R];
					grad[k][c] += k_ssim * (o[0] + 2.f * u * o[1] + v * o[2]);
				}
				__syncthreads();
			}
			if (a.w_s3im != 0.f) {
				// ---- S-B: vertical pass first: rows 0..41 (tile + 5), columns 0..35 (image x = tx0 - 2 + xx) ------------------
				for (int i = tid; i < LOSS_MID * LOSS_S3_VW; i += LOSS_THREADS) {
					const int yy = i / LOSS_S3_VW, xx = i % LOSS_S3_VW;
					float h1 = 0.f, h2 = 0.f, h11 = 0.f, h22 = 0.f, h12 = 0.f;
#pragma unroll
					for (int k = 0; k < LOSS_WIN; k++) {
						const int o = (yy + k) * LOSS_IN + xx + 2 * LOSS_R - 2;
						const float u = s_in1[o], v = s_in2[o], w = a.g[k];
						const float wu = w * u, wv = w * v;
						h1 += wu; h2 += wv; h11 = fmaf(wu, u, h11); h22 = fmaf(wv, v, h22); h12 = fmaf(wu, v, h12);
					}
					s_h[i] = h1; s_h[LOSS_MID * LOSS_S3_VW + i] = h2; s_h[2 * LOSS_MID * LOSS_S3_VW + i] = h11;
					s_h[3 * LOSS_MID * LOSS_S3_VW + i] = h22; s_h[4 * LOSS_MID * LOSS_S3_VW + i] = h12;
				}
				__syncthreads();
				// ---- S-C: per map point (rows 0..41, columns 0..33, image x = tx0 - 1 + xx) the ten replicas' maps; a point q
				// weighs pixel q-1 by ta, q by tb, q+1 by tc, so it keeps three sums per derivative -- what it owes to its
				// left neighbour, to itself and to its right neighbour -- in registers: no barrier inside the replica loop.
				float e[LOSS_S3_PTS][9];
#pragma unroll
				for (int k = 0; k < LOSS_S3_PTS; k++)
#pragma unroll
					for (int q = 0; q < 9; q++) e[k][q] = 0.f;
#pragma unroll
				for (int k = 0; k < LOSS_S3_PTS; k++) {
					const int i = tid + k * LOSS_THREADS;
					if (i >= LOSS_MID * LOSS_S3_W) continue;
					const int yy = i / LOSS_S3_W, xx = i % LOSS_S3_W;
					const int qx = tx0 - 1 + xx, qy = ty0 - LOSS_R + yy;
					if (!in_rect(qx, qy)) continue;
					float v[5][3];
#pragma unroll
					for (int q = 0; q < 5; q++)
#pragma unroll
						for (int t = 0; t < 3; t++) v[q][t] = s_h[q * LOSS_MID * LOSS_S3_VW + yy * LOSS_S3_VW + xx + t];
					const bool own = xx >= 1 && xx <= LOSS_T && yy >= LOSS_R && yy < LOSS_R + LOSS_T;
					for (int r = 0; r < LOSS_S3_REP; r++) {
						const float ta = a.s3[r][0], tb = a.s3[r][1], tc = a.s3[r][2];
						float m[5], dA, dB, dC;
#pragma unroll
						for (int q = 0; q < 5; q++) m[q] = ta * v[q][0] + tb * v[q][1] + tc * v[q][2];
						const float map = ssim_point(m[0], m[1], m[2], m[3], m[4], dA, dB, dC);
						if (own) sum_s3 += map;
						e[k][0] = fmaf(ta, dA, e[k][0]); e[k][1] = fmaf(tb, dA, e[k][1]); e[k][2] = fmaf(tc, dA, e[k][2]);
						e[k][3] = fmaf(ta, dB, e[k][3]); e[k][4] = fmaf(tb, dB, e[k][4]); e[k][5] = fmaf(tc, dB, e[k][5]);
						e[k][6] = fmaf(ta, dC, e[k][6]); e[k][7] = fmaf(tb, dC, e[k][7]); e[k][8] = fmaf(tc, dC, e[k][8]);
					}
				}
				__syncthreads();                    // the vertical-pass maps are dead: the nine sum maps take their place
#pragma unroll
				for (int k = 0; k < LOSS_S3_PTS; k++) {
					const int i = tid + k * LOSS_THREADS;
					if (i < LOSS_MID * LOSS_S3_W) {
#pragma unroll
						for (int q = 0; q < 9; q++) s_h[q * LOSS_MID * LOSS_S3_W + i] = e[k][q];
					}
				}
				__syncthreads();
#pragma unroll
				for (int k = 0; k < LOSS_PIX; k++) {
					const int p = tid + k * LOSS_THREADS, py = p / LOSS_T, px = p % LOSS_T;
					if (!in_rect(tx0 + px, ty0 + py)) continue;
					float o[3] = {0.f, 0.f, 0.f};
#pragma unroll
					for (int t = 0; t < LOSS_WIN; t++) {
						const float w = a.g[t];
						// pixel p (map column px + 1) collects "to the left" from q = p+1, "itself" from q = p, "to the right" from q = p-1
						const float* row = s_h + (py + t) * LOSS_S3_W + px;
#pragma unroll
						for (int q = 0; q < 3; q++) {
							const float* mq = row + 3 * q * LOSS_MID * LOSS_S3_W;
							o[q] = fmaf(w, mq[2] + mq[LOSS_MID * LOSS_S3_W + 1] + mq[2 * LOSS_MID * LOSS_S3_W], o[q]);
						}
					}
					const float u = s_in1[(py + 2 * LOSS_R) * LOSS_IN + px + 2 * LOSS_R], v = s_in2[(py + 2 * LOSS_R) * LOSS_IN + px + 2 * LOSS_R];
					grad[k][c] += k_s3 * (o[0] + 2.f * u * o[1] + v * o[2]);
				}
				__syncthreads();
			}
		}
	}

	// ---- masked L1 / L2 terms and the gradient rows of this thread's pixels ------------------------------------------------
	const float k_l1 = a.w_l1 / (3.f * (float)count), k_l2 = 2.f * a.w_mask / (float)count;
#pragma unroll
	for (int k = 0; k < LOSS_PIX; k++) {
		const int p = tid + k * LOSS_THREADS, x = tx0 + p % LOSS_T, y = ty0 + p / LOSS_T;
		if (x < 0 || y < 0 || x >= W || y >= H) continue;
		const size_t o = (size_t)y * W + x;
		const bool bound = a.bound ? __ldg(a.bound + o) == 1.f : true;
		float d_alpha = 0.f;
		if (bound) {
#pragma unroll
			for (int c = 0; c < NCH; c++) {
				const int ch = ch_first + c;
				const float d = __ldg(a.image + ch * N + o) - __ldg(a.gt + ch * N + o);
				sum_l1 += fabsf(d);
				grad[k][c] += d > 0.f ? k_l1 : (d < 0.f ? -k_l1 : 0.f);
			}
			if (ch_first == 0 && a.alpha && a.bkgd) {
				const float d = __ldg(a.alpha + o) - __ldg(a.bkgd + o);
				sum_l2 = fmaf(d, d, sum_l2);
				d_alpha = k_l2 * d;
			}
		}
		if (a.dL_dimage) {
#pragma unroll
			for (int c = 0; c < NCH; c++) a.dL_dimage[(ch_first + c) * N + o] = grad[k][c];
		}
		if (ch_first == 0 && a.dL_dalpha) a.dL_dalpha[o] = d_alpha;
	}

	// ---- block sums, then the last block adds the per-block sums in block order ----------------------------------------------
	float v[4] = {sum_l1, sum_l2, sum_ssim, sum_s3};
#pragma unroll
	for (int q = 0; q < 4; q++) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
		if (lane == 0) s_red[warp][q] = v[q];
	}
	__syncthreads();
	const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x, nblocks = gridDim.x * gridDim.y * gridDim.z;
	if (tid < 4) {
		float s = 0.f;
#pragma unroll
		for (int w = 0; w < LOSS_THREADS / 32; w++) s += s_red[w][tid];
		a.partial[4 * bid + tid] = s;
	}
	__threadfence();
	__syncthreads();
	if (tid == 0) s_last = atomicAdd(a.ticket, 1u) == (unsigned)nblocks - 1u;
	__syncthreads();
	if (!s_last) return;
	__threadfence();
	// a quarter of the block per loss: thread t adds blocks t, t+G, ... then a fixed-order tree over the G sums
	{
		constexpr int G = LOSS_THREADS / 4;
		const int q = tid / G, t = tid % G;
		float s = 0.f;
		for (int b0 = t; b0 < nblocks; b0 += G * 8) {
			float x[8];
#pragma unroll
			for (int i = 0; i < 8; i++) x[i] = b0 + G * i < nblocks ? __ldcg(a.partial + 4 * (size_t)(b0 + G * i) + q) : 0.f;
#pragma unroll
			for (int i = 0; i < 8; i++) s += x[i];
		}
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
		if (lane == 0) s_red[warp][0] = s;
	}
	__syncthreads();
	if (tid == 0) {
		*a.ticket = 0u;
		constexpr int WG = LOSS_THREADS / 32 / 4;       // warps per loss
		float S[4];
#pragma unroll
		for (int q = 0; q < 4; q++) {
			S[q] = 0.f;
#pragma unroll
			for (int w = 0; w < WG; w++) S[q] += s_red[q * WG + w][0];
		}
		const float S1 = S[0], S2 = S[1], S3 = S[2], S4 = S[3];
		const float crop = 3.f * (float)rw * (float)rh;
		const float l1 = S1 / (3.f * (float)count), l2 = (a.alpha && a.bkgd) ? S2 / (float)count : 0.f;
		const float ssim = S3 / crop, s3im = 1.f - S4 / (crop * (float)LOSS_S3_REP);
		a.losses[0] = l1; a.losses[1] = l2; a.losses[2] = ssim; a.losses[3] = s3im;
		float total = 0.f;
		if (a.w_l1 != 0.f) total += a.w_l1 * l1;
		if (a.w_mask != 0.f) total += a.w_mask * l2;
		if (a.w_ssim != 0.f) total += a.w_ssim * (1.f - ssim);
		if (a.w_s3im != 0.f) total += a.w_s3im * s3im;
		a.losses[4] = total;
	}
}

// Bounding rectangle of the non-zero pixels of a mask as cv2.boundingRect(mask.astype(np.uint8)) returns it
// (train_ZJU.py:115; (0,0,0,0) for an empty mask) and the number of pixels equal to 1 (the size of the boolean gathers at
// :111-112).  out int[8]: x, y, w, h, count, then scratch; the caller presets out[0..1] = 0x7f7f7f7f, out[2..3] =
// 0x80808080, out[4..7] = 0 (three memsets).  Integer atomics only: reproducible.
__global__ void __launch_bounds__(256)
mask_bounds_kernel(int W, int H, const float* __restrict__ mask, int* __restrict__ out)
{
	__shared__ bool s_last;
	int minx = 0x7f7f7f7f, miny = 0x7f7f7f7f, maxx = (int)0x80808080, maxy = (int)0x80808080, count = 0;
	const size_t N = (size_t)W * H;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (size_t)gridDim.x * blockDim.x) {
		const float v = __ldg(mask + i);
		count += v == 1.f;
		if ((unsigned char)(int)v != 0) {
			const int x = (int)(i % W), y = (int)(i / W);
			minx = min(minx, x); maxx = max(maxx, x); miny = min(miny, y); maxy = max(maxy, y);
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		minx = min(minx, __shfl_xor_sync(0xffffffffu, minx, o)); miny = min(miny, __shfl_xor_sync(0xffffffffu, miny, o));
		maxx = max(maxx, __shfl_xor_sync(0xffffffffu, maxx, o)); maxy = max(maxy, __shfl_xor_sync(0xffffffffu, maxy, o));
		count += __shfl_xor_sync(0xffffffffu, count, o);
	}
	if ((threadIdx.x & 31) == 0) {
		if (minx <= maxx) { atomicMin(out, minx); atomicMin(out + 1, miny); atomicMax(out + 2, maxx); atomicMax(out + 3, maxy); }
		if (count) atomicAdd(out + 4, count);
	}
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) s_last = atomicAdd((unsigned*)out + 5, 1u) == gridDim.x - 1u;
	__syncthreads();
	if (!s_last || threadIdx.x != 0) return;
	__threadfence();
	const int x0 = __ldcg(out), y0 = __ldcg(out + 1), x1 = __ldcg(out + 2), y1 = __ldcg(out + 3);
	if (x0 > x1) { out[0] = 0; out[1] = 0; out[2] = 0; out[3] = 0; }
	else { out[2] = x1 - x0 + 1; out[3] = y1 - y0 + 1; }
	out[5] = 0;
}

// dst *= *scale unless *scale == 1 (the upstream gradient of a loss that is the last node of the graph)
__global__ void __launch_bounds__(256)
scale_inplace_kernel(const float* __restrict__ scale, float* __restrict__ a, size_t na, float* __restrict__ b, size_t nb)
{
	const float s = __ldg(scale);
	if (s == 1.f) return;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < na + nb; i += (size_t)gridDim.x * blockDim.x) {
		if (i < na) a[i] *= s; else b[i - na] *= s;
	}
}

static_assert(9 * LOSS_MID * LOSS_S3_W <= LOSS_SM_H + LOSS_SM_D, "the nine S3IM sum maps fit the pass + derivative buffers");

} // namespace moss

