// Densification bookkeeping right behind the backward (first half).
//
// The reference updates three per-Gaussian statistics after every iteration with boolean-mask indexing
// (train_ZJU.py:171-174, scene/gaussian_model.py:816-818):
//     visibility_filter = radii > 0                                                  (gaussian_renderer/__init__.py:128)
//     max_radii2D[vis]        = max(max_radii2D[vis], radii[vis])
//     xyz_gradient_accum[vis] += norm(viewspace_point_tensor.grad[vis, :2], dim=-1, keepdim=True)
//     denom[vis]              += 1
// i.e. three `nonzero` host synchronisations and ~10 gather/scatter launches.  Here: one elementwise pass, no sync.
// 24 B read + up to 12 B read-modify-written per Gaussian: HBM-bound.
#pragma once
#include "common.cuh"

namespace moss {

__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ dL_dmean2D,
	float* __restrict__ xyz_gradient_accum, float* __restrict__ denom, float* __restrict__ max_radii2D)
{
	const int i = blockIdx.x * 256 + threadIdx.x;
	if (i >= P) return;
	const int r = radii[i];
	if (r <= 0) return;
	if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
	if (xyz_gradient_accum) {
		const float gx = dL_dmean2D[3 * i], gy = dL_dmean2D[3 * i + 1];
		xyz_gradient_accum[i] += __fsqrt_rn(__fmaf_rn(gy, gy, __fmul_rn(gx, gx)));
	}
	if (denom) denom[i] += 1.0f;
}

} // namespace moss
