// Shared definitions for the sm_100a splat rasterizer kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace moss {

constexpr int TILE_X = 16;          // DGR/cuda_rasterizer/config.h:16-17 (tile = 16x16 px)
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
constexpr int NCH = 3;              // DGR/cuda_rasterizer/config.h:15

// ---------------------------------------------------------------------------------------------
// Explicitly-rounded fp32 helpers.  The integer outputs (radii, tile rects, depth key bits) must be
// bit-identical to the reference, which depends on WHERE nvcc fuses mul+add into fma in the
// reference's expressions.  Intrinsics are never re-contracted, so the rounding sequence below is
// pinned independent of compiler mood.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float frcp(float a) { return __frcp_rn(a); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }

// m[12+i] + (z*m[8+i] + (x*m[i] + y*m[4+i]))   -- rounding order of the reference's
// transformPoint4x3/4x4 expressions (DGR/cuda_rasterizer/auxiliary.h:58-77) as nvcc contracts them.
__device__ __forceinline__ float affine_row(float x, float y, float z, float m0, float m4, float m8, float m12)
{
	return fadd(m12, ffma(z, m8, ffma(x, m0, fmul(y, m4))));
}
// a2*b2 + (a0*b0 + a1*b1)   -- GLM mat3*mat3 element (glm/detail/type_mat3x3.inl:509-517) after contraction
__device__ __forceinline__ float dot3_glm(float a0, float b0, float a1, float b1, float a2, float b2)
{
	return ffma(a2, b2, ffma(a0, b0, fmul(a1, b1)));
}

__host__ __device__ __forceinline__ size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---------------------------------------------------------------------------------------------
// Workspace carving (caller-owned memory; layout private to this library).
// ---------------------------------------------------------------------------------------------
constexpr int TILE_SUB = 8;        // sub-counters per tile (ImageWS::tile_count)
constexpr int GRAD_ACC_VEC = 3;    // float4 slots per Gaussian in GeomWS::grad_acc (48 B)
constexpr int GRAD_EXT_VEC = 2;    // + float4 slots in GeomWS::grad_ext, touched by MOSS_FLAG_DETERMINISTIC only
constexpr int MAX_VIEWS = 32;      // views per batched call (MOSS_MAX_VIEWS); the per-view camera table travels as a kernel argument
constexpr int VIEW_SYNC_WORDS = 64; // look-back words of the per-view tile scan, carved behind tile_count (one memset clears both)

// A batched call renders V views of the SAME Gaussians in one launch sequence.  Everything per view is laid out
// view-major: "virtual Gaussian" id' = v*P + i, "virtual tile" t' = v*T + t, pixel v*N + p.  V = 1 is the plain call.

struct GeomWS {
	float*    depths;         // [P]   view-space z (sort key low bits)
	float2*   means2D;        // [P]   pixel-space centre
	float4*   conic_opacity;  // [P]   (conic.x, conic.y, conic.z, opacity)
	float4*   feat;           // [P]   (r, g, b, depth): one 16 B gather per blended instance
	float*    cov3D;          // [6P]  only written when built from scale/rotation
	uint32_t* tiles_touched;  // [P]
	ushort4*  rect;           // [P]   tile rect (xmin, ymin, xmax, ymax)
	uint8_t*  clamped;        // [P]   bit c set <=> SH colour channel c was clamped at 0
	float4*   grad_acc;       // [3P]  blend-backward accumulators, 48 B per Gaussian, VECTOR atomics:
	                          //       [0] = (dL_dcolor rgb, dL_dopacity)  [1] = (dL_dmean2D xy, dL_dconic xy)
	                          //       [2] = (dL_dconic w, -, -, -).  Rows of VISIBLE Gaussians are zeroed by the forward
	                          //       preprocess, consumed and re-zeroed by the backward preprocess (rows of culled
	                          //       Gaussians are never touched): no memset launch, no bytes for culled Gaussians.
	float4*   grad_ext;       // [2P]  MOSS_FLAG_DETERMINISTIC only: the mode views grad_acc (48 B) + grad_ext (32 B) as
	                          //       nine int64 fixed-point sums (colour rgb, opacity, mean2D xy | conic x y w):
	                          //       integer adds commute, so the result does not depend on the order in which blocks
	                          //       arrive.  Cleared by a memset in the deterministic backward; untouched otherwise.
};
// slot s (0..8) of a Gaussian's deterministic accumulator
__device__ __forceinline__ unsigned long long* det_slot(const GeomWS& g, size_t id, int s)
{
	return s < 6 ? reinterpret_cast<unsigned long long*>(g.grad_acc + GRAD_ACC_VEC * id) + s
	             : reinterpret_cast<unsigned long long*>(g.grad_ext + GRAD_EXT_VEC * id) + (s - 6);
}
struct ImageWS {
	float*    final_T;        // [N]
	uint32_t* n_contrib;      // [N]
	uint32_t* tile_count;     // [T][TILE_SUB] (+ VIEW_SYNC_WORDS look-back words)  per-tile instance counts, split over TILE_SUB sub-counters picked by the
	                          //      counting warp (a busy tile's single counter is a serialisation point in L2: ~1700
	                          //      same-address atomics took ~9 us at ZJU shape); after the scan: the scatter cursors
	uint32_t* tile_total;     // [T]  instances per tile
	uint2*    ranges;         // [T]  [first, last) into the sorted list
	uint32_t* tile_order;     // [T]  launch order of the blend kernels: tiles by descending instance count
	uint32_t* counters;       // [8]  num_rendered, overflow, capacity, non-empty tiles, backward segments, checkpoint records,
	                          //      bits of max|upstream gradient| (deterministic backward), reserved
	uint32_t* view_busy;      // [MAX_VIEWS]    non-empty tiles of every view
	uint32_t* view_base;      // [MAX_VIEWS+1]  first instance of every view in the sorted list (+ total)
	uint32_t* view_sync;      // = tile_count + T*TILE_SUB: (instances of view v) + 1 once block v of the scan knows it
};
struct BinWS {
	unsigned long long* keys; // [cap] (depth_bits << 32 | gaussian id), grouped per tile, sorted in place
	uint32_t* point_list;     // [cap] sorted gaussian ids
	unsigned long long* keys_sorted; // [cap] (tile << 32 | depth_bits), only with EXPORT_KEYS
	unsigned long long* keys_tmp;    // [cap] bucketed copy of a tile's keys: tiles deeper than the shared-memory sort holds
};

template <typename T>
__host__ __device__ inline T* carve(char*& p, size_t count)
{
	p = (char*)align_up((size_t)p, 256);
	T* r = (T*)p;
	p += count * sizeof(T);
	return r;
}
__host__ __device__ inline GeomWS carve_geom(char* base, size_t P, size_t* bytes = nullptr)
{
	char* p = base;
	GeomWS g;
	g.depths = carve<float>(p, P);
	g.means2D = carve<float2>(p, P);
	g.conic_opacity = carve<float4>(p, P);
	g.feat = carve<float4>(p, P);
	g.cov3D = carve<float>(p, 6 * P);
	g.tiles_touched = carve<uint32_t>(p, P);
	g.rect = carve<ushort4>(p, P);
	g.clamped = carve<uint8_t>(p, P);
	g.grad_acc = carve<float4>(p, GRAD_ACC_VEC * P);
	g.grad_ext = carve<float4>(p, GRAD_EXT_VEC * P);
	if (bytes) *bytes = (size_t)(p - base) + 256;
	return g;
}
__host__ __device__ inline ImageWS carve_image(char* base, size_t N, size_t T, size_t* bytes = nullptr)
{
	char* p = base;
	ImageWS s;
	s.final_T = carve<float>(p, N);
	s.n_contrib = carve<uint32_t>(p, N);
	s.tile_count = carve<uint32_t>(p, T * TILE_SUB + VIEW_SYNC_WORDS);
	s.view_sync = s.tile_count + T * TILE_SUB;
	s.tile_total = carve<uint32_t>(p, T);
	s.ranges = carve<uint2>(p, T);
	s.tile_order = carve<uint32_t>(p, T);
	s.counters = carve<uint32_t>(p, 8);
	s.view_busy = carve<uint32_t>(p, MAX_VIEWS);
	s.view_base = carve<uint32_t>(p, MAX_VIEWS + 1);
	if (bytes) *bytes = (size_t)(p - base) + 256;
	return s;
}
__host__ __device__ inline BinWS carve_bin(char* base, size_t cap, bool export_keys, size_t* bytes = nullptr)
{
	char* p = base;
	BinWS b;
	b.keys = carve<unsigned long long>(p, cap);
	b.point_list = carve<uint32_t>(p, cap);
	b.keys_sorted = export_keys ? carve<unsigned long long>(p, cap) : nullptr;
	b.keys_tmp = carve<unsigned long long>(p, cap);
	if (bytes) *bytes = (size_t)(p - base) + 256;
	return b;
}

// MOSS_FLAG_DETERMINISTIC: fixed-point scale of the blend-backward accumulators, a power of two derived from the
// largest |upstream gradient| m of the call (bits in counters[6]).  A Gaussian's sums are bounded by about
// 2^17 (largest per-pixel factor: 0.5*W*|d|*conic) x 2^12 pixels x m, so S = 2^(31 - ceil(log2 m)) keeps S*sum below
// 2^62 while resolving 2^-31 m, finer than fp32 at m.
__device__ __forceinline__ float det_scale_from_bits(uint32_t bits, bool inverse)
{
	const float m = __uint_as_float(bits);
	int e = 0;
	if (m > 0.f && m < 3.0e38f) frexpf(m, &e);
	return ldexpf(1.0f, inverse ? (e - 31) : (31 - e));
}

// Optional checkpoint workspace (moss_ckpt_workspace_bytes): lets the blend backward split a deep tile's instance
// list into segments of CKPT_SEG that run as independent blocks.  The forward stores every pixel's running state
// (T, C rgb, D) at the segment boundaries and its final sums; a backward segment starts from the boundary state
// instead of walking everything behind it (blend.cuh).
constexpr int CKPT_SEG = 128;     // = the blend kernels' staging batch for 2- and 4-warp blocks (whole-tile blocks stage 256: no checkpoints)
constexpr int CKPT_REC_FLOATS = 5 * TILE_PIX;   // one boundary record of one tile: [component][pixel slot]
struct CkptWS {
	float4*   fin;        // [N]  final (C0, C1, C2, D) of every pixel, without background
	uint32_t* seg_base;   // [T]  first boundary record of the tile
	uint2*    seg_map;    // [max_segs]  (tile, segment) of backward block group k, deep tiles first (roughly)
	float*    rec;        // [max_recs][5][256]
	uint32_t  max_segs, max_recs;
};
__host__ __device__ inline CkptWS carve_ckpt(char* base, size_t N, size_t T, size_t cap, size_t* bytes = nullptr)
{
	char* p = base;
	CkptWS c;
	c.max_recs = (uint32_t)(cap / CKPT_SEG);
	c.max_segs = (uint32_t)(cap / CKPT_SEG + T);
	c.fin = carve<float4>(p, N);
	c.seg_base = carve<uint32_t>(p, T);
	c.seg_map = carve<uint2>(p, c.max_segs);
	c.rec = carve<float>(p, (size_t)c.max_recs * CKPT_REC_FLOATS);
	if (bytes) *bytes = (size_t)(p - base) + 256;
	return c;
}
__host__ __device__ inline CkptWS no_ckpt()
{
	CkptWS c; c.fin = nullptr; c.seg_base = nullptr; c.seg_map = nullptr; c.rec = nullptr; c.max_segs = 0; c.max_recs = 0;
	return c;
}

// Camera table of a (batched) call.  The matrices live in device memory (they are torch tensors in the reference
// API: DGR/diff_gaussian_rasterization/__init__.py:160-172); each block stages its view's in shared memory once.
// Passed by value as a __grid_constant__ kernel argument, so a batched call needs no device-side table and
// stays CUDA-graph capturable.
struct ViewCam {
	const float* view;    // [16]
	const float* proj;    // [16]
	const float* campos;  // [3]
	const float* bg;      // [3]
	float tanfovx, tanfovy, focal_x, focal_y;
};
struct ViewBatch {
	int V;                // views in this call (1 = the reference's single-view call)
	int P;                // Gaussians (shared by all views)
	int W, H, grid_x, grid_y;
	int T;                // tiles per view
	int D, M;
	float scale_modifier;
	ViewCam cam[MAX_VIEWS];
};

} // namespace moss
