// Fused AdamW step over the Gaussian parameter groups (second half).
//
// The reference builds `torch.optim.AdamW(l, lr=0.0, eps=1e-15)` over six per-Gaussian tensors with their own learning
// rates (scene/gaussian_model.py:203-225) and calls `optimizer.step()` once per iteration (train_ZJU.py:189): with
// torch's default implementation that is ~10 elementwise launches per group.  Here ONE launch updates up to 8 groups:
// every block takes a 1024-element chunk of one group (block -> group through the table of chunk offsets).
//
// Arithmetic per element, in torch's order (torch/optim/adamw.py, _single_tensor_adamw; amsgrad off, maximize off):
//     p  *= 1 - lr * weight_decay
//     m   = m + (g - m) * (1 - beta1)                        (Tensor.lerp_)
//     v   = v * beta2 + (1 - beta2) * g * g                  (mul_ + addcmul_)
//     p  -= (lr / (1 - beta1^t)) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)
// 16 B read + 12 B written per element: HBM-bound.
#pragma once
#include "common.cuh"

namespace moss {

constexpr int ADAM_MAX_GROUPS = 8;
constexpr int ADAM_CHUNK = 1024;          // elements per block (256 threads x float4)

struct AdamGroups {
	float* param[ADAM_MAX_GROUPS];
	const float* grad[ADAM_MAX_GROUPS];
	float* exp_avg[ADAM_MAX_GROUPS];
	float* exp_avg_sq[ADAM_MAX_GROUPS];
	unsigned long long numel[ADAM_MAX_GROUPS];
	unsigned chunk_end[ADAM_MAX_GROUPS];  // exclusive end of the group's chunks in the grid
	float decay[ADAM_MAX_GROUPS];         // 1 - lr * weight_decay
	float step_size[ADAM_MAX_GROUPS];     // lr / (1 - beta1^t)
	float bc2_sqrt[ADAM_MAX_GROUPS];      // sqrt(1 - beta2^t)   (applied as a division, like torch)
	int n;
	float w1, beta2, w2, eps;             // 1 - beta1, beta2, 1 - beta2, eps
};

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, float decay, float w1, float beta2,
	float w2, float step_size, float bc2_sqrt, float eps)
{
	p = __fmul_rn(p, decay);
	m = __fmaf_rn(__fsub_rn(g, m), w1, m);                       // lerp for weight < 0.5: start + weight * (end - start)
	v = __fmaf_rn(__fmul_rn(w2, g), g, __fmul_rn(v, beta2));     // addcmul_: self + value * t1 * t2
	const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps);
	p = __fmaf_rn(-step_size, __fdiv_rn(m, denom), p);           // addcdiv_: self + value * (t1 / t2)
}

__global__ void __launch_bounds__(256)
adamw_kernel(AdamGroups a)
{
	int gi = 0;
#pragma unroll
	for (int k = 0; k < ADAM_MAX_GROUPS - 1; k++) gi += (k < a.n - 1 && blockIdx.x >= a.chunk_end[k]) ? 1 : 0;
	const unsigned first = gi ? a.chunk_end[gi - 1] : 0u;
	const unsigned long long base = (unsigned long long)(blockIdx.x - first) * ADAM_CHUNK + 4ull * threadIdx.x;
	const unsigned long long n = a.numel[gi];
	if (base >= n) return;
	float* P = a.param[gi] + base;
	const float* G = a.grad[gi] + base;
	float* M = a.exp_avg[gi] + base;
	float* V = a.exp_avg_sq[gi] + base;
	const float decay = a.decay[gi], ss = a.step_size[gi], bc2 = a.bc2_sqrt[gi];
	const bool vec = base + 4 <= n && ((((uintptr_t)P | (uintptr_t)G | (uintptr_t)M | (uintptr_t)V) & 15) == 0);
	if (vec) {
		float4 p = *reinterpret_cast<float4*>(P), m = *reinterpret_cast<float4*>(M), v = *reinterpret_cast<float4*>(V);
		const float4 g = *reinterpret_cast<const float4*>(G);
		adamw_one(p.x, g.x, m.x, v.x, decay, a.w1, a.beta2, a.w2, ss, bc2, a.eps);
		adamw_one(p.y, g.y, m.y, v.y, decay, a.w1, a.beta2, a.w2, ss, bc2, a.eps);
		adamw_one(p.z, g.z, m.z, v.z, decay, a.w1, a.beta2, a.w2, ss, bc2, a.eps);
		adamw_one(p.w, g.w, m.w, v.w, decay, a.w1, a.beta2, a.w2, ss, bc2, a.eps);
		*reinterpret_cast<float4*>(P) = p; *reinterpret_cast<float4*>(M) = m; *reinterpret_cast<float4*>(V) = v;
	} else {
		for (int k = 0; k < 4 && base + k < n; k++) {
			float p = P[k], m = M[k], v = V[k];
			adamw_one(p, G[k], m, v, decay, a.w1, a.beta2, a.w2, ss, bc2, a.eps);
			P[k] = p; M[k] = m; V[k] = v;
		}
	}
}

} // namespace moss
