// C-ABI entry points (include/moss_splat.h) and launch orchestration for the sm_100a splat rasterizer.
//
// Forward launch sequence (5 launches, zero host synchronisation in ASYNC mode):
//   memset(tile_count) -> K1 preprocess+count -> K2 tile scan -> K3 scatter -> K4 sort+blend
// versus the reference's preprocess -> CUB scan -> blocking D2H -> duplicateWithKeys -> CUB radix sort
// (~6 passes) -> memset -> identifyTileRanges -> render (DGR/cuda_rasterizer/rasterizer_impl.cu:198-341).
// In synchronous mode the host waits only for an event recorded right after K2 (so K3/K4 are already queued
// behind it and the GPU never idles) to learn num_rendered and detect binning-buffer overflow.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>      // header-only (v3): resolves the tool's injection library at the first call, a no-op without one
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <mutex>
#include <atomic>
#include <vector>
#include <algorithm>
#include "../../include/moss_splat.h"
#include "common.cuh"
#include "preprocess.cuh"
#include "preprocess_bwd.cuh"
#include "binning.cuh"
#include "blend.cuh"
#include "skin.cuh"
#include "nearest.cuh"
#include "densify.cuh"
#include "adamw.cuh"
#include "peer_reduce.cuh"
#include "push_exchange.cuh"
#include "smpl_frame.cuh"
#include "loss.cuh"

using namespace moss;

namespace {

thread_local int g_last_cuda_error = 0;

#define MOSS_CUDA_TRY(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { g_last_cuda_error = (int)e__; return MOSS_ERR_CUDA; } } while (0)

inline int check_launch(cudaStream_t s, unsigned flags)
{
	cudaError_t e = cudaGetLastError();
	if (e == cudaSuccess && (flags & MOSS_FLAG_DEBUG)) e = cudaStreamSynchronize(s);
	if (e != cudaSuccess) { g_last_cuda_error = (int)e; return MOSS_ERR_CUDA; }
	return MOSS_OK;
}
#define MOSS_CHECK_LAUNCH(s, flags) do { int r__ = check_launch(s, flags); if (r__ != MOSS_OK) return r__; } while (0)

// ---- instrumentation: launch counter + optional per-stage CUDA-event timing ----------------------------
std::atomic<unsigned long long> g_launches{0};
struct Profiler {
	static constexpr int CAP = 8192;
	struct Rec { cudaEvent_t a, b; int stage, dev; };
	bool on = false;
	std::vector<Rec> rec;              // events belong to the device that was current when they were created
	int n = 0;
	double ms[MOSS_STAGE_COUNT] = {0};
	unsigned long long cnt[MOSS_STAGE_COUNT] = {0};
	std::mutex mu;
	int drain()
	{
		int cur = 0;
		cudaGetDevice(&cur);
		int status = MOSS_OK;
		for (int i = 0; i < n; i++) {
			float t = 0.f;
			if (rec[i].dev != cur) cudaSetDevice(rec[i].dev);
			if (cudaEventSynchronize(rec[i].b) != cudaSuccess || cudaEventElapsedTime(&t, rec[i].a, rec[i].b) != cudaSuccess) status = MOSS_ERR_CUDA;
			else { ms[rec[i].stage] += t; cnt[rec[i].stage]++; }
			if (rec[i].dev != cur) cudaSetDevice(cur);
		}
		n = 0;
		return status;
	}
	void begin(int st, cudaStream_t s)
	{
		std::lock_guard<std::mutex> lk(mu);
		if (n == CAP) drain();
		int cur = 0;
		cudaGetDevice(&cur);
		if ((int)rec.size() <= n) {
			Rec r; r.stage = st; r.dev = cur;
			cudaEventCreate(&r.a); cudaEventCreate(&r.b);
			rec.push_back(r);
		} else if (rec[n].dev != cur) {         // a pooled pair made on another device: replace it
			cudaEventDestroy(rec[n].a); cudaEventDestroy(rec[n].b);
			cudaEventCreate(&rec[n].a); cudaEventCreate(&rec[n].b);
			rec[n].dev = cur;
		}
		rec[n].stage = st;
		cudaEventRecord(rec[n].a, s);
	}
	void end(cudaStream_t s)
	{
		std::lock_guard<std::mutex> lk(mu);
		cudaEventRecord(rec[n].b, s);
		n++;
	}
};
Profiler g_prof;

// NVTX ranges: one range per C-ABI call ("moss_rasterize_forward", ...) and one per kernel launch
// inside it, named like the profiler's stages, so `ncu --nvtx --nvtx-include "moss_rasterize_backward/"` (or an nsys
// timeline) attributes every launch.  Off by default; MOSS_NVTX=1 in the environment or moss_nvtx_enable(1) turns it on.
std::atomic<int> g_nvtx{-1};          // -1: not decided yet (read MOSS_NVTX once)
inline bool nvtx_on()
{
	int v = g_nvtx.load(std::memory_order_relaxed);
	if (v < 0) {
		const char* e = getenv("MOSS_NVTX");
		v = (e && *e && *e != '0') ? 1 : 0;
		g_nvtx.store(v, std::memory_order_relaxed);
	}
	return v != 0;
}
const char* stage_name(int stage)
{
	static const char* names[MOSS_STAGE_COUNT] = {"preprocess_forward", "tile_scan", "scatter_keys", "blend_forward",
		"blend_backward", "preprocess_backward", "other", "tile_sort", "skin_forward", "skin_backward", "peer_allreduce", "nearest_vertex", "densify_stats", "adamw_step", "sum_views", "push_gather", "smpl_frame_forward", "smpl_frame_backward", "image_losses"};
	return (stage >= 0 && stage < MOSS_STAGE_COUNT) ? names[stage] : "?";
}
struct ApiRange {
	bool on;
	explicit ApiRange(const char* name) : on(nvtx_on()) { if (on) nvtxRangePushA(name); }
	~ApiRange() { if (on) nvtxRangePop(); }
};
#define MOSS_API_RANGE(name) ApiRange api_range__(name)

struct StageScope {
	cudaStream_t s; bool on, nvtx;
	StageScope(int stage, cudaStream_t s_) : s(s_), on(g_prof.on), nvtx(nvtx_on())
	{
		g_launches.fetch_add(1, std::memory_order_relaxed);
		if (nvtx) nvtxRangePushA(stage_name(stage));
		if (on) g_prof.begin(stage, s);
	}
	~StageScope() { if (on) g_prof.end(s); if (nvtx) nvtxRangePop(); }
};
#define MOSS_STAGE(stage, stream) StageScope stage_scope__(stage, stream)

// Pinned, device-mapped slots through which K2 hands num_rendered to the host without a blocking copy.  One pool per
// device (an event may only be recorded on a stream of the device it was created on); a slot stays marked busy from
// acquire() until its forward call has read it, so concurrent callers never share one.
constexpr int MAX_DEVICES = 64;
struct HostSlots {
	static constexpr int N = 64;
	int* host = nullptr;       // cudaHostAlloc'ed, mapped
	cudaEvent_t ev[N];
	bool busy[N] = {false};
	int next = 0;
	bool ok = false;
	std::mutex mu;
	int acquire()              // -1: pool exhausted (64 forwards in flight on this device) or CUDA error
	{
		std::lock_guard<std::mutex> lk(mu);
		if (!ok) {
			if (cudaHostAlloc((void**)&host, N * sizeof(int), cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) return -1;
			for (int i = 0; i < N; i++)
				if (cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) != cudaSuccess) return -1;
			ok = true;
		}
		for (int k = 0; k < N; k++) {
			const int s = (next + k) % N;
			if (!busy[s]) { busy[s] = true; next = (s + 1) % N; return s; }
		}
		return -1;
	}
	void release(int s)
	{
		std::lock_guard<std::mutex> lk(mu);
		if (s >= 0 && s < N) busy[s] = false;
	}
};
HostSlots g_slots_dev[MAX_DEVICES];
inline HostSlots* device_slots()
{
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return nullptr;
	return &g_slots_dev[dev];
}
struct SlotLease {             // releases the slot on every return path of the forward
	HostSlots* pool = nullptr; int slot = -1;
	~SlotLease() { if (pool && slot >= 0) pool->release(slot); }
};

// cudaFuncSetAttribute is per device: remember where the sort kernel's dynamic shared memory was already raised
std::atomic<bool> g_sort_attr_set[MAX_DEVICES];
std::atomic<bool> g_smpl_attr_set[MAX_DEVICES];
std::atomic<bool> g_loss_attr_set[MAX_DEVICES];

static inline bool skin_valid(const MossSkin* k)
{
	if (!k || k->P < 0 || k->V <= 0) return false;
	if (k->P == 0) return true;
	if (!k->xyz || !k->vert_ids || !k->weights || !k->A_big || !k->A_tgt || !k->Rg || !k->Th) return false;
	if ((k->scaling != nullptr) != (k->rotation != nullptr)) return false;
	if (((uintptr_t)k->weights & 15) || ((uintptr_t)k->lbs_delta & 15)) return false;
	return true;
}
static inline SkinArgs make_skin(const MossSkin* k)
{
	SkinArgs a;
	a.P = k->P; a.xyz = k->xyz; a.vert_ids = k->vert_ids; a.weights = k->weights; a.lbs_delta = k->lbs_delta;
	a.A_big = k->A_big; a.A_tgt = k->A_tgt; a.poseoff_big = k->poseoff_big; a.shapeoff = k->shapeoff;
	a.poseoff_tgt = k->poseoff_tgt; a.Rg = k->Rg; a.Th = k->Th; a.scaling = k->scaling; a.rotation = k->rotation;
	a.scale_modifier = k->scale_modifier;
	return a;
}
inline bool scene_valid(const MossScene* sc, bool own_camera = true, const MossSkin* skin = nullptr)
{
	if (!sc || sc->P < 0 || sc->W <= 0 || sc->H <= 0) return false;
	if (own_camera && !sc->bg) return false;
	if (sc->P == 0) return true;
	if (!sc->opacities) return false;
	if (own_camera && (!sc->viewmatrix || !sc->projmatrix)) return false;
	if ((sc->shs != nullptr) == (sc->colors_precomp != nullptr)) return false;
	if (skin) {
		// fused skin + preprocess: position and covariance come from the skinning inputs
		if (sc->means3D || sc->scales || sc->rotations || sc->cov3D_precomp) return false;
		if (!skin_valid(skin) || skin->P != sc->P || !skin->scaling) return false;
	} else {
		if (!sc->means3D) return false;
		const bool sr = (sc->scales != nullptr) && (sc->rotations != nullptr);
		if (sr == (sc->cov3D_precomp != nullptr)) return false;
		if (!sr && (sc->scales || sc->rotations)) return false;
	}
	if (sc->shs && ((own_camera && !sc->campos) || sc->D < 0 || sc->D > 3 || sc->M < (sc->D + 1) * (sc->D + 1))) return false;
	if (((sc->W + TILE_X - 1) / TILE_X) > 65535 || ((sc->H + TILE_Y - 1) / TILE_Y) > 65535) return false;
	return true;
}

// Camera table of a call.  `views` == nullptr: the single camera described by the scene itself.
inline bool make_batch(const MossScene* sc, int V, const MossView* views, ViewBatch* vb)
{
	if (V < 1 || V > MAX_VIEWS) return false;
	vb->V = V; vb->P = sc->P;
	vb->W = sc->W; vb->H = sc->H;
	vb->grid_x = (sc->W + TILE_X - 1) / TILE_X; vb->grid_y = (sc->H + TILE_Y - 1) / TILE_Y;
	vb->T = vb->grid_x * vb->grid_y;
	vb->D = sc->D; vb->M = sc->M;
	vb->scale_modifier = sc->scale_modifier;
	for (int v = 0; v < V; v++) {
		ViewCam& c = vb->cam[v];
		float tx, ty;
		if (views) {
			c.view = views[v].viewmatrix; c.proj = views[v].projmatrix; c.campos = views[v].campos; c.bg = views[v].bg;
			tx = views[v].tanfovx; ty = views[v].tanfovy;
		} else {
			c.view = sc->viewmatrix; c.proj = sc->projmatrix; c.campos = sc->campos; c.bg = sc->bg;
			tx = sc->tanfovx; ty = sc->tanfovy;
		}
		if (!c.view || !c.proj || !c.bg || (sc->shs && !c.campos)) return false;
		c.tanfovx = tx; c.tanfovy = ty;
		c.focal_y = sc->H / (2.0f * ty);   // DGR rasterizer_impl.cu:224-225
		c.focal_x = sc->W / (2.0f * tx);
	}
	return true;
}

// Blend block shape (1 whole tile / 2 half / 3 quarter).  Small images have few busy tiles and the deepest one sets
// the launch time, so its patches are spread over more SMs; large images have enough tiles to fill the machine
// and prefer fewer, larger blocks (less duplicated staging).  A batch of views of a body-sized image has enough busy
// tiles but they stay deep (1000-4000 instances): half-tile blocks measured best there (tools/batch_split_probe.py:
// 23 ZJU views 1251 us vs 1351 us with whole tiles, 1264 us with quarters).
inline int blend_split(unsigned flags, int T, int V)
{
	const int f = (int)((flags & MOSS_FLAG_SPLIT_MASK) >> MOSS_FLAG_SPLIT_SHIFT);
	if (f) return f;
	if (V > 1) return T <= 6144 ? 2 : 1;
	return T <= 2048 ? 3 : (T <= 6144 ? 2 : 1);
}

// The optional checkpoint workspace is used when it is large enough for this call's binning capacity.
inline CkptWS scene_ckpt(const MossScene* sc, uint32_t capacity, size_t N, int T, int V, unsigned flags)
{
	// whole-tile (8-warp) blocks stage 256 instances per round and would skip every other boundary
	if (blend_split(flags, T, V) == 1 && CKPT_SEG < 256) return no_ckpt();
	if (!sc->ckpt_ws || sc->ckpt_bytes < moss_ckpt_workspace_bytes_batch((long long)capacity, sc->W, sc->H, V)) return no_ckpt();
	return carve_ckpt((char*)sc->ckpt_ws, N, (size_t)T * V, (size_t)capacity);
}

__global__ void reset_overflow_kernel(uint32_t* counters, uint32_t capacity)
{
	counters[1] = counters[0] > capacity ? 1u : 0u;
	counters[2] = capacity;
	counters[4] = 0u; counters[5] = 0u;
}

// grid = (ceil(N / 256), V)
__global__ void fill_background_kernel(const __grid_constant__ ViewBatch vb, int N, float* __restrict__ color,
	float* __restrict__ depth, float* __restrict__ alpha)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= N) return;
	const size_t v = blockIdx.y;
	const float* bg = vb.cam[v].bg;
	float* c = color + v * 3 * N;
	c[i] = bg[0]; c[N + i] = bg[1]; c[2 * N + i] = bg[2];
	depth[v * N + i] = 0.f; alpha[v * N + i] = 0.f;
}

} // namespace

extern "C" {

int moss_abi_version(void) { return MOSS_ABI_VERSION; }
size_t moss_sizeof_scene(void) { return sizeof(MossScene); }
unsigned long long moss_kernel_launches(void) { return g_launches.load(); }
int moss_profile_enable(int on)
{
	std::lock_guard<std::mutex> lk(g_prof.mu);
	g_prof.on = on != 0;
	return MOSS_OK;
}
int moss_profile_read(double* ms, unsigned long long* launches)
{
	std::lock_guard<std::mutex> lk(g_prof.mu);
	const int r = g_prof.drain();
	for (int i = 0; i < MOSS_STAGE_COUNT; i++) {
		if (ms) ms[i] = g_prof.ms[i];
		if (launches) launches[i] = g_prof.cnt[i];
		g_prof.ms[i] = 0; g_prof.cnt[i] = 0;
	}
	return r;
}
const char* moss_stage_name(int stage) { return stage_name(stage); }
int moss_nvtx_enable(int on)
{
	g_nvtx.store(on ? 1 : 0, std::memory_order_relaxed);
	return MOSS_OK;
}
int moss_last_cuda_error(void) { return g_last_cuda_error; }

const char* moss_error_string(int status)
{
	switch (status) {
	case MOSS_OK: return "ok";
	case MOSS_ERR_BINNING_OVERFLOW: return "binning workspace too small for num_rendered (re-run with MOSS_FLAG_REBIN)";
	case MOSS_ERR_INVALID_ARGUMENT: return "invalid argument";
	case MOSS_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
	case MOSS_ERR_CUDA: return "CUDA runtime error";
	case MOSS_ERR_STATIC_SCRATCH_LIMIT: return "image exceeds the static scratch of moss_image_losses_static (at most 4096 tiles of 32x32 pixels, e.g. 2048 x 2048); use moss_image_losses with a caller buffer";
	default: return "unknown status";
	}
}

size_t moss_geom_workspace_bytes_batch(int P, int V)
{
	size_t b = 0;
	carve_geom(nullptr, (size_t)(P > 0 ? P : 0) * (size_t)(V > 0 ? V : 1), &b);
	return b;
}
size_t moss_geom_workspace_bytes(int P) { return moss_geom_workspace_bytes_batch(P, 1); }
size_t moss_image_workspace_bytes_batch(int W, int H, int V)
{
	size_t b = 0;
	const size_t v = (size_t)(V > 0 ? V : 1);
	const size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
	carve_image(nullptr, (size_t)W * H * v, T * v, &b);
	return b;
}
size_t moss_image_workspace_bytes(int W, int H) { return moss_image_workspace_bytes_batch(W, H, 1); }
size_t moss_binning_workspace_bytes_ex(long long capacity, int export_keys)
{
	size_t b = 0;
	carve_bin(nullptr, (size_t)(capacity > 0 ? capacity : 0), export_keys != 0, &b);
	return b;
}
size_t moss_binning_workspace_bytes(long long capacity) { return moss_binning_workspace_bytes_ex(capacity, 0); }
size_t moss_ckpt_workspace_bytes_batch(long long capacity, int W, int H, int V)
{
	size_t b = 0;
	const size_t v = (size_t)(V > 0 ? V : 1);
	const size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
	carve_ckpt(nullptr, (size_t)W * H * v, T * v, (size_t)(capacity > 0 ? capacity : 0), &b);
	return b;
}
size_t moss_ckpt_workspace_bytes(long long capacity, int W, int H) { return moss_ckpt_workspace_bytes_batch(capacity, W, H, 1); }
long long moss_binning_capacity(size_t bytes, int export_keys)
{
	const size_t per = export_keys ? 28 : 20;
	long long lo = 0, hi = (long long)(bytes / per) + 1;   // largest cap with bytes_ex(cap) <= bytes
	while (lo < hi) {
		const long long mid = lo + (hi - lo + 1) / 2;
		if (moss_binning_workspace_bytes_ex(mid, export_keys) <= bytes) lo = mid; else hi = mid - 1;
	}
	return lo;
}

namespace {

// per-view row of the batched backward's scratch area (floats; every segment 16 B-aligned)
struct BatchRow { size_t opacity, color, mean3D, cov3D, sh, scale, rot, total; };
inline BatchRow batch_row(size_t P, size_t M)
{
	auto pad = [](size_t n) { return (n + 3) / 4 * 4; };
	BatchRow r; size_t o = 0;
	r.opacity = o; o += pad(P);
	r.color = o; o += pad(3 * P);
	r.mean3D = o; o += pad(3 * P);
	r.cov3D = o; o += pad(6 * P);
	r.sh = o; o += pad(3 * M * P);
	r.scale = o; o += pad(3 * P);
	r.rot = o; o += pad(4 * P);
	r.total = o;
	return r;
}

// validates a MossPush and derives the producer-side routing table, the gather kernel's arguments and its grid
int make_gather(const MossPush* push, PushCfg* pc, GatherArgs* ga, unsigned* gather_blocks)
{
	if (push->world < 1 || push->world > PEER_MAX_WORLD || push->rank < 0 || push->rank >= push->world || !push->grad_base ||
		((uintptr_t)push->grad_base & 15) || (push->numel & 3) || push->numel == 0 || push->numel > 0xfffffff0ull ||
		!push->grad_ptrs || !push->stage_ptrs || !push->signal_ptrs || !push->error_word_dev ||
		push->signal_pad_bytes < moss_peer_signal_bytes(push->world))
		return MOSS_ERR_INVALID_ARGUMENT;
	const size_t S = moss_push_slice_floats(push->numel, push->world);
	pc->base = push->grad_base; pc->numel = (uint32_t)push->numel; pc->S = (uint32_t)S;
	memset(ga, 0, sizeof(*ga));
	for (int o = 0; o < push->world; o++) {
		if (!push->grad_ptrs[o] || !push->stage_ptrs[o] || !push->signal_ptrs[o]) return MOSS_ERR_INVALID_ARGUMENT;
		pc->stage_mine[o] = (float*)push->stage_ptrs[o] + (size_t)push->rank * S;
		ga->stage_mine[o] = pc->stage_mine[o];
		ga->grad[o] = (float*)push->grad_ptrs[o];
		ga->pads[o] = (uint32_t*)push->signal_ptrs[o];
	}
	ga->stage = (const float*)push->stage_ptrs[push->rank];
	ga->mc_grad = (float*)push->multicast_grad;
	ga->err = push->error_word_dev;
	ga->numel = pc->numel; ga->S = pc->S; ga->rank = push->rank; ga->world = push->world;
	// every rank derives the same grid from (numel, world, pad size): the flag slots are indexed by it
	size_t blocks = (S / 4 + (size_t)GATHER_THREADS - 1) / (size_t)GATHER_THREADS;      // one float4 per thread and owner, if the pads allow
	const size_t pad_blocks = push->signal_pad_bytes / (sizeof(uint32_t) * 2 * (size_t)push->world);
	blocks = std::min(blocks, std::min(pad_blocks, (size_t)PEER_MAX_BLOCKS));
	*gather_blocks = (unsigned)std::max<size_t>(blocks, 1);
	return MOSS_OK;
}

int forward_impl(cudaStream_t stream, const MossScene* sc, int V, const MossView* views, unsigned flags,
	float* out_color, float* out_depth, float* out_alpha, int* radii, int* num_rendered,
	const MossSkin* skin = nullptr, float* sk_means3D = nullptr, float* sk_transforms = nullptr)
{
	if (!scene_valid(sc, views == nullptr, skin) || !out_color || !out_depth || !out_alpha) return MOSS_ERR_INVALID_ARGUMENT;
	if (skin && V != 1) return MOSS_ERR_INVALID_ARGUMENT;
	if (sc->P > 0 && !radii) return MOSS_ERR_INVALID_ARGUMENT;
	if ((unsigned long long)sc->P * (unsigned long long)V >= 0xffffffffull) return MOSS_ERR_INVALID_ARGUMENT;   // id' = v*P + i is 32 bits
	if (num_rendered) *num_rendered = (flags & MOSS_FLAG_ASYNC) ? -1 : 0;
	ViewBatch vb;
	if (!make_batch(sc, V, views, &vb)) return MOSS_ERR_INVALID_ARGUMENT;
	const int P = sc->P, W = sc->W, H = sc->H, N = W * H;
	const int T = vb.T, TV = T * V;
	const size_t NV = (size_t)N * V;
	if (sc->image_bytes < moss_image_workspace_bytes_batch(W, H, V) || !sc->image_ws) return MOSS_ERR_WORKSPACE_TOO_SMALL;
	ImageWS im = carve_image((char*)sc->image_ws, NV, (size_t)TV);
	if (P == 0) {
		// DGR/rasterize_points.cu:83: nothing rendered; outputs are background / zeros, state says "no contributors"
		{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		fill_background_kernel<<<dim3((N + 255) / 256, V), 256, 0, stream>>>(vb, N, out_color, out_depth, out_alpha); }
		MOSS_CUDA_TRY(cudaMemsetAsync(sc->image_ws, 0, moss_image_workspace_bytes_batch(W, H, V) - 256, stream));
		MOSS_CHECK_LAUNCH(stream, flags);
		return MOSS_OK;
	}
	if (sc->geom_bytes < moss_geom_workspace_bytes_batch(P, V) || !sc->geom_ws) return MOSS_ERR_WORKSPACE_TOO_SMALL;
	const bool export_keys = (flags & MOSS_FLAG_EXPORT_KEYS) != 0;
	const long long cap_ll = sc->binning_ws ? moss_binning_capacity(sc->binning_bytes, export_keys) : 0;
	const uint32_t capacity = (uint32_t)(cap_ll > 0xfffffff0ll ? 0xfffffff0ll : cap_ll);
	GeomWS g = carve_geom((char*)sc->geom_ws, (size_t)P * V);
	BinWS b = carve_bin((char*)sc->binning_ws, capacity, export_keys);
	const CkptWS ck = scene_ckpt(sc, capacity, NV, T, V, flags);

	// synchronous mode: K2 drops the instance count into a pinned mapped slot and the host waits for an event recorded
	// right behind K2 (K3/K4 are queued behind it already); without a free slot it reads the counters with a copy
	SlotLease lease;
	if (!(flags & (MOSS_FLAG_ASYNC | MOSS_FLAG_REBIN))) {
		lease.pool = device_slots();
		lease.slot = lease.pool ? lease.pool->acquire() : -1;
	}
	const int slot = lease.slot;
	const dim3 pre_grid((P + PRE_THREADS - 1) / PRE_THREADS, V);

	if (!(flags & MOSS_FLAG_REBIN)) {
		MOSS_CUDA_TRY(cudaMemsetAsync(im.tile_count, 0, ((size_t)TV * TILE_SUB + VIEW_SYNC_WORDS) * sizeof(uint32_t), stream));
		{ MOSS_STAGE(MOSS_STAGE_PREPROCESS, stream);
		SkinFused sk;
		memset(&sk, 0, sizeof(sk));
		if (skin) {
			sk.a = make_skin(skin); sk.means3D = sk_means3D; sk.transforms = sk_transforms;
			preprocess_forward_kernel<true><<<pre_grid, PRE_THREADS, 0, stream>>>(
				vb, nullptr, sc->shs, sc->colors_precomp, sc->opacities, nullptr, nullptr, nullptr, sc->prefiltered, g, im.tile_count, radii, sk);
		} else {
			preprocess_forward_kernel<false><<<pre_grid, PRE_THREADS, 0, stream>>>(
				vb, sc->means3D, sc->shs, sc->colors_precomp, sc->opacities, sc->scales, sc->rotations,
				sc->cov3D_precomp, sc->prefiltered, g, im.tile_count, radii, sk);
		} }
		MOSS_CHECK_LAUNCH(stream, flags);
		int* host_R = nullptr;
		if (slot >= 0) MOSS_CUDA_TRY(cudaHostGetDevicePointer((void**)&host_R, lease.pool->host + slot, 0));
		{ MOSS_STAGE(MOSS_STAGE_TILE_SCAN, stream);
		tile_scan_kernel<<<V, SCAN_THREADS, 0, stream>>>(T, V, im.tile_count, im.tile_total, im.ranges, im.counters, capacity, host_R,
			im.tile_order, im.view_busy, im.view_base, im.view_sync); }
		MOSS_CHECK_LAUNCH(stream, flags);
		if (slot >= 0) MOSS_CUDA_TRY(cudaEventRecord(lease.pool->ev[slot], stream));
	} else {
		{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		reset_overflow_kernel<<<1, 1, 0, stream>>>(im.counters, capacity); }
		MOSS_CHECK_LAUNCH(stream, flags);
	}

	{ MOSS_STAGE(MOSS_STAGE_SCATTER, stream);
	scatter_keys_kernel<<<dim3((P + SCATTER_THREADS - 1) / SCATTER_THREADS, V), SCATTER_THREADS, 0, stream>>>(P, vb.grid_x, T, radii, g,
		im.ranges, im.tile_count, im.counters, b.keys); }
	MOSS_CHECK_LAUNCH(stream, flags);
	{ MOSS_STAGE(MOSS_STAGE_TILE_SORT, stream);
	// opt in to > 48 KB of dynamic shared memory (the attribute is per device)
	int cur_dev = 0;
	MOSS_CUDA_TRY(cudaGetDevice(&cur_dev));
	if (cur_dev >= 0 && cur_dev < MAX_DEVICES && !g_sort_attr_set[cur_dev].load(std::memory_order_acquire)) {
		MOSS_CUDA_TRY(cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SortSmem)));
		g_sort_attr_set[cur_dev].store(true, std::memory_order_release);
	}
	tile_sort_kernel<<<TV, TILE_PIX, sizeof(SortSmem), stream>>>(im, b, ck, T, V); }
	MOSS_CHECK_LAUNCH(stream, flags);
	{ MOSS_STAGE(MOSS_STAGE_BLEND_FWD, stream);
	const bool cull = !(flags & MOSS_FLAG_NO_CULL);
	switch (blend_split(flags, T, V)) {
	case 1:
		if (cull) blend_forward_kernel<true, 8><<<TV, 256, 0, stream>>>(vb, g, im, b, ck, out_color, out_depth, out_alpha);
		else blend_forward_kernel<false, 8><<<TV, 256, 0, stream>>>(vb, g, im, b, ck, out_color, out_depth, out_alpha);
		break;
	case 2:
		if (cull) blend_forward_kernel<true, 4><<<2 * TV, 128, 0, stream>>>(vb, g, im, b, ck, out_color, out_depth, out_alpha);
		else blend_forward_kernel<false, 4><<<2 * TV, 128, 0, stream>>>(vb, g, im, b, ck, out_color, out_depth, out_alpha);
		break;
	default:
		if (cull) blend_forward_kernel<true, 2><<<4 * TV, 64, 0, stream>>>(vb, g, im, b, ck, out_color, out_depth, out_alpha);
		else blend_forward_kernel<false, 2><<<4 * TV, 64, 0, stream>>>(vb, g, im, b, ck, out_color, out_depth, out_alpha);
		break;
	} }
	MOSS_CHECK_LAUNCH(stream, flags);

	if (flags & MOSS_FLAG_ASYNC) return MOSS_OK;
	int R;
	if (slot >= 0) {
		MOSS_CUDA_TRY(cudaEventSynchronize(lease.pool->ev[slot]));
		R = *(volatile int*)(lease.pool->host + slot);
	} else {
		uint32_t c[4];
		MOSS_CUDA_TRY(cudaMemcpyAsync(c, im.counters, sizeof(c), cudaMemcpyDeviceToHost, stream));
		MOSS_CUDA_TRY(cudaStreamSynchronize(stream));
		R = (int)c[0];
	}
	if (num_rendered) *num_rendered = R;
	if ((uint32_t)R > capacity) return MOSS_ERR_BINNING_OVERFLOW;
	return MOSS_OK;
}

int backward_impl(cudaStream_t stream, const MossScene* sc, int V, const MossView* views, unsigned flags, const int* radii,
	const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, void* tmp_ws, size_t tmp_bytes,
	const MossPush* push = nullptr, const MossSkin* skin = nullptr, const SkinGradOut* skin_out = nullptr)
{
	if (!scene_valid(sc, views == nullptr, skin)) return MOSS_ERR_INVALID_ARGUMENT;
	if (skin && (V != 1 || push || !skin_out)) return MOSS_ERR_INVALID_ARGUMENT;
	PushCfg pc = no_push();
	GatherArgs ga;
	unsigned gather_blocks = 0;
	if (push) {
		const int r = make_gather(push, &pc, &ga, &gather_blocks);
		if (r != MOSS_OK) return r;
	}
	const int P = sc->P, W = sc->W, H = sc->H, N = W * H;
	if (P == 0) return MOSS_OK;
	if (!radii || !dL_dpix || !dL_dmean2D || !dL_dopacity || !dL_dcolor ||
		(!skin && (!dL_dmean3D || !dL_dcov3D)) || (sc->shs && !dL_dsh) || (sc->scales && (!dL_dscale || !dL_drot)))
		return MOSS_ERR_INVALID_ARGUMENT;
	if (skin) {
		if (!skin_out->d_xyz || !skin_out->d_scaling || !skin_out->d_rotation || (skin_out->d_lbs_delta && !skin->lbs_delta) ||
			((uintptr_t)skin_out->d_rotation & 15) || ((uintptr_t)skin_out->d_lbs_delta & 15)) return MOSS_ERR_INVALID_ARGUMENT;
		if (skin_out->d_A_tgt) MOSS_CUDA_TRY(cudaMemsetAsync(skin_out->d_A_tgt, 0, sizeof(float) * NJ * 16, stream));
		if (skin_out->d_poseoff_tgt) MOSS_CUDA_TRY(cudaMemsetAsync(skin_out->d_poseoff_tgt, 0, sizeof(float) * 3 * (size_t)skin->V, stream));
	}
	ViewBatch vb;
	if (!make_batch(sc, V, views, &vb)) return MOSS_ERR_INVALID_ARGUMENT;
	const int T = vb.T, TV = T * V;
	const size_t NV = (size_t)N * V;
	if (sc->image_bytes < moss_image_workspace_bytes_batch(W, H, V) || sc->geom_bytes < moss_geom_workspace_bytes_batch(P, V))
		return MOSS_ERR_WORKSPACE_TOO_SMALL;
	if (V > 1 && (!tmp_ws || tmp_bytes < moss_batch_grad_workspace_bytes(P, sc->M, V))) return MOSS_ERR_WORKSPACE_TOO_SMALL;
	const bool export_keys = (flags & MOSS_FLAG_EXPORT_KEYS) != 0;
	const long long cap_ll = sc->binning_ws ? moss_binning_capacity(sc->binning_bytes, export_keys) : 0;
	const uint32_t capacity = (uint32_t)(cap_ll > 0xfffffff0ll ? 0xfffffff0ll : cap_ll);
	GeomWS g = carve_geom((char*)sc->geom_ws, (size_t)P * V);
	ImageWS im = carve_image((char*)sc->image_ws, NV, (size_t)TV);
	BinWS b = carve_bin((char*)sc->binning_ws, capacity, export_keys);
	const CkptWS ck = scene_ckpt(sc, capacity, NV, T, V, flags);
	// work units: tiles, or (with checkpoints) at most capacity/CKPT_SEG + T segments; surplus blocks exit at once
	const unsigned units = ck.rec ? ck.max_segs : (unsigned)TV;
	const int det = (flags & MOSS_FLAG_DETERMINISTIC) ? 1 : 0;
	if (det) {
		MOSS_CUDA_TRY(cudaMemsetAsync(im.counters + 6, 0, sizeof(uint32_t), stream));
		MOSS_CUDA_TRY(cudaMemsetAsync(g.grad_ext, 0, sizeof(float4) * GRAD_EXT_VEC * (size_t)P * V, stream));
		{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		upstream_maxabs_kernel<<<(int)std::min<size_t>((NV + 255) / 256, 592), 256, 0, stream>>>((int)NV, dL_dpix, dL_ddepth, dL_dalpha, im.counters + 6); }
		MOSS_CHECK_LAUNCH(stream, flags);
	}

	{ MOSS_STAGE(MOSS_STAGE_BLEND_BWD, stream);
	switch (blend_split(flags, T, V)) {
	case 1: blend_backward_kernel<8><<<units, 256, 0, stream>>>(vb, g, im, b, ck, dL_dpix, dL_ddepth, dL_dalpha, det); break;
	case 2: blend_backward_kernel<4><<<2 * units, 128, 0, stream>>>(vb, g, im, b, ck, dL_dpix, dL_ddepth, dL_dalpha, det); break;
	default: blend_backward_kernel<2><<<4 * units, 64, 0, stream>>>(vb, g, im, b, ck, dL_dpix, dL_ddepth, dL_dalpha, det); break;
	} }
	MOSS_CHECK_LAUNCH(stream, flags);

	ParamGradOut po;
	BatchRow row = batch_row((size_t)P, (size_t)sc->M);
	if (V == 1) {
		po.color = dL_dcolor; po.opacity = dL_dopacity; po.mean3D = dL_dmean3D; po.cov3D = dL_dcov3D;
		po.sh = sc->shs ? dL_dsh : nullptr; po.scale = sc->scales ? dL_dscale : nullptr; po.rot = sc->scales ? dL_drot : nullptr;
		po.view_stride = 0;
	} else {
		float* t = (float*)tmp_ws;
		po.color = t + row.color; po.opacity = t + row.opacity; po.mean3D = t + row.mean3D; po.cov3D = t + row.cov3D;
		po.sh = sc->shs ? t + row.sh : nullptr; po.scale = sc->scales ? t + row.scale : nullptr; po.rot = sc->scales ? t + row.rot : nullptr;
		po.view_stride = row.total;
	}
	{ MOSS_STAGE(MOSS_STAGE_PREPROCESS_BWD, stream);
	SkinFused sk;
	SkinGradOut sg;
	memset(&sk, 0, sizeof(sk)); memset(&sg, 0, sizeof(sg));
	if (skin) {
		sk.a = make_skin(skin); sg = *skin_out;
		preprocess_backward_kernel<true><<<dim3((P + PRE_THREADS - 1) / PRE_THREADS, V), PRE_THREADS, 0, stream>>>(
			vb, nullptr, radii, sc->shs, nullptr, nullptr, nullptr, g, dL_dmean2D, dL_dconic, po, det, im.counters, no_push(), sk, sg);
	} else {
		preprocess_backward_kernel<false><<<dim3((P + PRE_THREADS - 1) / PRE_THREADS, V), PRE_THREADS, 0, stream>>>(
			vb, sc->means3D, radii, sc->shs, sc->scales, sc->rotations, sc->cov3D_precomp, g,
			dL_dmean2D, dL_dconic, po, det, im.counters, no_push(), sk, sg);
	} }
	MOSS_CHECK_LAUNCH(stream, flags);
	if (V > 1) {
		SumViewsArgs a;
		memset(&a, 0, sizeof(a));
		a.V = V; a.view_stride = row.total;
		const size_t p = (size_t)P, m = (size_t)sc->M;
		float* dst[7] = {dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, sc->shs ? dL_dsh : nullptr, sc->scales ? dL_dscale : nullptr, sc->scales ? dL_drot : nullptr};
		const float* src[7] = {po.opacity, po.color, po.mean3D, po.cov3D, po.sh, po.scale, po.rot};
		const size_t n[7] = {p, 3 * p, 3 * p, 6 * p, 3 * m * p, 3 * p, 4 * p};
		size_t biggest = 0;
		for (int k = 0; k < 7; k++) { a.dst[k] = dst[k]; a.src[k] = src[k]; a.n[k] = dst[k] ? n[k] : 0; biggest = std::max(biggest, a.n[k]); }
		const unsigned gx = (unsigned)std::min<size_t>((biggest / 4 + SUMV_THREADS - 1) / SUMV_THREADS + 1, 148 * 8);
		{ MOSS_STAGE(MOSS_STAGE_SUM_VIEWS, stream);
		sum_views_kernel<<<dim3(gx, 7), SUMV_THREADS, 0, stream>>>(a, pc); }
		MOSS_CHECK_LAUNCH(stream, flags);
	}
	if (push) {
		// one view: the rows were written locally (K7's small strided stores make poor NVLink packets) and the gather
		// kernel opens with a coalesced push; a batch: the view sum already pushed them
		ga.push_src = V == 1 ? push->grad_base : nullptr;
		{ MOSS_STAGE(MOSS_STAGE_PUSH_GATHER, stream);
		push_gather_kernel<<<gather_blocks, GATHER_THREADS, 0, stream>>>(ga); }
		MOSS_CHECK_LAUNCH(stream, flags);
	}
	return MOSS_OK;
}

} // namespace

size_t moss_sizeof_view(void) { return sizeof(MossView); }
size_t moss_sizeof_push(void) { return sizeof(MossPush); }
size_t moss_push_slice_floats(size_t numel, int world)
{
	const size_t w = (size_t)(world > 0 ? world : 1);
	return ((numel + w - 1) / w + 3) / 4 * 4;
}
int moss_push_allreduce(void* stream_, const MossPush* push)
{
	MOSS_API_RANGE("moss_push_allreduce");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!push) return MOSS_ERR_INVALID_ARGUMENT;
	PushCfg pc; GatherArgs ga; unsigned blocks = 0;
	const int r = make_gather(push, &pc, &ga, &blocks);
	if (r != MOSS_OK) return r;
	ga.push_src = push->grad_base;
	{ MOSS_STAGE(MOSS_STAGE_PUSH_GATHER, stream);
	push_gather_kernel<<<blocks, GATHER_THREADS, 0, stream>>>(ga); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}
int moss_rasterize_backward_push(void* stream, const MossScene* sc, int V, const MossView* views, unsigned flags, const int* radii,
	const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, void* tmp_ws, size_t tmp_bytes,
	const MossPush* push)
{
	MOSS_API_RANGE("moss_rasterize_backward_push");
	if (!push || (!views && V != 1)) return MOSS_ERR_INVALID_ARGUMENT;
	return backward_impl((cudaStream_t)stream, sc, V, views, flags, radii, dL_dpix, dL_ddepth, dL_dalpha, dL_dmean2D, dL_dconic,
		dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, tmp_ws, tmp_bytes, push);
}
size_t moss_batch_grad_workspace_bytes(int P, int M, int V)
{
	if (V <= 1 || P <= 0) return 0;
	return batch_row((size_t)P, (size_t)(M > 0 ? M : 0)).total * sizeof(float) * (size_t)V;
}

int moss_rasterize_forward(void* stream, const MossScene* sc, unsigned flags,
	float* out_color, float* out_depth, float* out_alpha, int* radii, int* num_rendered)
{
	MOSS_API_RANGE("moss_rasterize_forward");
	return forward_impl((cudaStream_t)stream, sc, 1, nullptr, flags, out_color, out_depth, out_alpha, radii, num_rendered);
}
int moss_skin_rasterize_forward(void* stream, const MossScene* sc, const MossSkin* skin, unsigned flags,
	float* out_color, float* out_depth, float* out_alpha, int* radii, int* num_rendered, float* means3D, float* transforms)
{
	MOSS_API_RANGE("moss_skin_rasterize_forward");
	if (!skin) return MOSS_ERR_INVALID_ARGUMENT;
	return forward_impl((cudaStream_t)stream, sc, 1, nullptr, flags, out_color, out_depth, out_alpha, radii, num_rendered, skin, means3D, transforms);
}
int moss_skin_rasterize_backward(void* stream, const MossScene* sc, const MossSkin* skin, unsigned flags, const int* radii,
	const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dsh,
	float* d_xyz, float* d_scaling, float* d_rotation, float* d_lbs_delta, float* d_A_tgt, float* d_poseoff_tgt)
{
	MOSS_API_RANGE("moss_skin_rasterize_backward");
	if (!skin) return MOSS_ERR_INVALID_ARGUMENT;
	SkinGradOut sg;
	sg.d_xyz = d_xyz; sg.d_scaling = d_scaling; sg.d_rotation = d_rotation; sg.d_lbs_delta = d_lbs_delta;
	sg.d_A_tgt = d_A_tgt; sg.d_poseoff_tgt = d_poseoff_tgt;
	return backward_impl((cudaStream_t)stream, sc, 1, nullptr, flags, radii, dL_dpix, dL_ddepth, dL_dalpha, dL_dmean2D, dL_dconic,
		dL_dopacity, dL_dcolor, nullptr, nullptr, dL_dsh, nullptr, nullptr, nullptr, 0, nullptr, skin, &sg);
}
int moss_rasterize_forward_batch(void* stream, const MossScene* sc, int V, const MossView* views, unsigned flags,
	float* out_color, float* out_depth, float* out_alpha, int* radii, int* num_rendered)
{
	MOSS_API_RANGE("moss_rasterize_forward_batch");
	if (!views) return MOSS_ERR_INVALID_ARGUMENT;
	return forward_impl((cudaStream_t)stream, sc, V, views, flags, out_color, out_depth, out_alpha, radii, num_rendered);
}
int moss_rasterize_backward(void* stream, const MossScene* sc, unsigned flags, const int* radii,
	const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot)
{
	MOSS_API_RANGE("moss_rasterize_backward");
	if (!dL_dconic) return MOSS_ERR_INVALID_ARGUMENT;
	return backward_impl((cudaStream_t)stream, sc, 1, nullptr, flags, radii, dL_dpix, dL_ddepth, dL_dalpha, dL_dmean2D, dL_dconic,
		dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, nullptr, 0);
}
int moss_rasterize_backward_batch(void* stream, const MossScene* sc, int V, const MossView* views, unsigned flags, const int* radii,
	const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, void* tmp_ws, size_t tmp_bytes)
{
	MOSS_API_RANGE("moss_rasterize_backward_batch");
	if (!views) return MOSS_ERR_INVALID_ARGUMENT;
	return backward_impl((cudaStream_t)stream, sc, V, views, flags, radii, dL_dpix, dL_ddepth, dL_dalpha, dL_dmean2D, dL_dconic,
		dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, tmp_ws, tmp_bytes);
}

int moss_mark_visible(void* stream_, int P, const float* means3D, const float* viewmatrix,
	const float* projmatrix, uint8_t* present)
{
	MOSS_API_RANGE("moss_mark_visible");
	(void)projmatrix;
	cudaStream_t stream = (cudaStream_t)stream_;
	if (P < 0) return MOSS_ERR_INVALID_ARGUMENT;
	if (P == 0) return MOSS_OK;
	if (!means3D || !viewmatrix || !present) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
	mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, viewmatrix, present); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

namespace {
__global__ void expand_clamped_kernel(int P, const uint8_t* __restrict__ bits, uint8_t* __restrict__ out)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= P) return;
	const uint8_t c = bits[i];
	out[3 * i] = c & 1; out[3 * i + 1] = (c >> 1) & 1; out[3 * i + 2] = (c >> 2) & 1;
}
__global__ void extract_rgb_kernel(int P, const float4* __restrict__ feat, float* __restrict__ out)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= P) return;
	const float4 f = feat[i];
	out[3 * i] = f.x; out[3 * i + 1] = f.y; out[3 * i + 2] = f.z;
}
}

long long moss_export_state_batch(void* stream_, const MossScene* sc, int V, int which, void* dst, long long max_bytes)
{
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!sc || !dst || V < 1 || V > MAX_VIEWS) return -MOSS_ERR_INVALID_ARGUMENT;
	const size_t P = (size_t)sc->P * V, N = (size_t)sc->W * sc->H * V;
	const int gx = (sc->W + TILE_X - 1) / TILE_X, gy = (sc->H + TILE_Y - 1) / TILE_Y;
	const size_t T = (size_t)gx * gy * V;
	GeomWS g = carve_geom((char*)sc->geom_ws, P);
	ImageWS im = carve_image((char*)sc->image_ws, N, T);
	uint32_t counters[4] = {0, 0, 0, 0};
	if (which == MOSS_STATE_KEYS_SORTED || which == MOSS_STATE_POINT_LIST) {
		if (cudaMemcpyAsync(counters, im.counters, sizeof(counters), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
			cudaStreamSynchronize(stream) != cudaSuccess) return -MOSS_ERR_CUDA;
	}
	const size_t R = counters[0];
	const void* src = nullptr;
	long long n = -1;
	switch (which) {
	case MOSS_STATE_DEPTHS: src = g.depths; n = 4 * P; break;
	case MOSS_STATE_MEANS2D: src = g.means2D; n = 8 * P; break;
	case MOSS_STATE_COV3D: src = g.cov3D; n = 24 * P; break;
	case MOSS_STATE_CONIC_OPACITY: src = g.conic_opacity; n = 16 * P; break;
	case MOSS_STATE_TILES_TOUCHED: src = g.tiles_touched; n = 4 * P; break;
	case MOSS_STATE_RANGES: src = im.ranges; n = 8 * T; break;
	case MOSS_STATE_N_CONTRIB: src = im.n_contrib; n = 4 * N; break;
	case MOSS_STATE_FINAL_T: src = im.final_T; n = 4 * N; break;
	case MOSS_STATE_COUNTERS: src = im.counters; n = 16; break;
	case MOSS_STATE_VIEW_BASE: src = im.view_base; n = 4 * ((long long)V + 1); break;
	case MOSS_STATE_CLAMPED:
		n = 3 * P;
		if (n > max_bytes) return -MOSS_ERR_WORKSPACE_TOO_SMALL;
		if (P) expand_clamped_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>((int)P, g.clamped, (uint8_t*)dst);
		return n;
	case MOSS_STATE_RGB:
		n = 12 * P;
		if (n > max_bytes) return -MOSS_ERR_WORKSPACE_TOO_SMALL;
		if (P) extract_rgb_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>((int)P, g.feat, (float*)dst);
		return n;
	case MOSS_STATE_KEYS_SORTED:
	case MOSS_STATE_POINT_LIST: {
		// counters[2] is the capacity the forward carved the binning workspace with; keys_sorted (present only
		// with MOSS_FLAG_EXPORT_KEYS) lies behind keys and point_list, so one carve resolves both.
		if (R > (size_t)counters[2]) return -MOSS_ERR_WORKSPACE_TOO_SMALL;
		BinWS b = carve_bin((char*)sc->binning_ws, (size_t)counters[2], true);
		if (which == MOSS_STATE_KEYS_SORTED) {
			if (moss_binning_workspace_bytes_ex(counters[2], 1) > sc->binning_bytes) return -MOSS_ERR_WORKSPACE_TOO_SMALL;
			src = b.keys_sorted; n = 8 * R;
		} else { src = b.point_list; n = 4 * R; }
		break;
	}
	default: return -MOSS_ERR_INVALID_ARGUMENT;
	}
	if (n > max_bytes) return -MOSS_ERR_WORKSPACE_TOO_SMALL;
	if (n > 0 && cudaMemcpyAsync(dst, src, (size_t)n, cudaMemcpyDeviceToDevice, stream) != cudaSuccess) return -MOSS_ERR_CUDA;
	return n;
}
long long moss_export_state(void* stream, const MossScene* sc, int which, void* dst, long long max_bytes)
{
	return moss_export_state_batch(stream, sc, 1, which, dst, max_bytes);
}

// ---- fused skinning + covariance front end ---------------------------------------------------------------
size_t moss_sizeof_skin(void) { return sizeof(MossSkin); }

int moss_skin_forward(void* stream_, const MossSkin* k, float* means3D, float* cov3D, float* transforms,
	float* bweights, float* translation)
{
	MOSS_API_RANGE("moss_skin_forward");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!skin_valid(k)) return MOSS_ERR_INVALID_ARGUMENT;
	if (k->P == 0) return MOSS_OK;
	if (!means3D || (cov3D && !k->scaling) || ((uintptr_t)bweights & 15)) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_SKIN, stream);
		skin_forward_kernel<<<(k->P + SKIN_THREADS - 1) / SKIN_THREADS, SKIN_THREADS, 0, stream>>>(
			make_skin(k), means3D, cov3D, transforms, bweights, translation); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_skin_backward(void* stream_, const MossSkin* k, const float* g_means3D, const float* g_cov3D,
	float* d_xyz, float* d_scaling, float* d_rotation, float* d_lbs_delta, float* d_A_tgt, float* d_poseoff_tgt)
{
	MOSS_API_RANGE("moss_skin_backward");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!skin_valid(k)) return MOSS_ERR_INVALID_ARGUMENT;
	if (d_A_tgt) MOSS_CUDA_TRY(cudaMemsetAsync(d_A_tgt, 0, sizeof(float) * NJ * 16, stream));
	if (d_poseoff_tgt) MOSS_CUDA_TRY(cudaMemsetAsync(d_poseoff_tgt, 0, sizeof(float) * 3 * (size_t)k->V, stream));
	if (k->P == 0) return MOSS_OK;
	if (!g_means3D || !d_xyz) return MOSS_ERR_INVALID_ARGUMENT;
	if (k->scaling && (!d_scaling || !d_rotation)) return MOSS_ERR_INVALID_ARGUMENT;
	if (d_lbs_delta && !k->lbs_delta) return MOSS_ERR_INVALID_ARGUMENT;
	if (((uintptr_t)d_rotation & 15) || ((uintptr_t)d_lbs_delta & 15)) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_SKIN_BWD, stream);
		skin_backward_kernel<<<(k->P + SKIN_THREADS - 1) / SKIN_THREADS, SKIN_THREADS, 0, stream>>>(
			make_skin(k), g_means3D, g_cov3D, d_xyz, d_scaling, d_rotation, d_lbs_delta, d_A_tgt, d_poseoff_tgt); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

// ---- per-frame SMPL tables -------------------------------------------------------------------------------------------
// scratch layout (backward only): [0,16) ticket (zero between calls), chain dL/dR [24*9 floats], partial sums [pose_blocks][208]
static int smpl_shape_blocks(int V) { return (V + SMPL_THREADS - 1) / SMPL_THREADS; }
static int smpl_pose_blocks(int V) { return (3 * V + SMPL_ROWS_PER_BLOCK - 1) / SMPL_ROWS_PER_BLOCK; }
static int smpl_bwd_blocks(int V) { return (3 * V + SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK - 1) / (SMPL_BWD_STAGES * SMPL_ROWS_PER_BLOCK); }
size_t moss_smpl_frame_scratch_bytes(int V)
{
	if (V < 1) return 16;
	return 16 + ((size_t)NJ * 9 + (size_t)smpl_bwd_blocks(V) * (NPF + 1)) * sizeof(float);
}
size_t moss_sizeof_smpl_frame(void) { return sizeof(MossSmplFrame); }
static bool smpl_frame_valid(const MossSmplFrame* f)
{
	return f && f->V >= 1 && f->n_shapes >= 0 && f->n_shapes <= f->shape_stride && f->parents && f->poses &&
		(f->shapes || f->n_shapes == 0);
}

int moss_smpl_joint_tables(void* stream_, int V, int S, const float* v_template, const float* shapedirs,
	const float* J_regressor, float* J_template, float* J_shapedirs)
{
	MOSS_API_RANGE("moss_smpl_joint_tables");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (V < 1 || S < 0 || !v_template || !J_regressor || !J_template || (S > 0 && (!shapedirs || !J_shapedirs)))
		return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		smpl_joint_tables_kernel<<<dim3(NJ, 1 + S), SMPL_THREADS, 0, stream>>>(V, S, v_template, shapedirs, J_regressor, J_template, J_shapedirs); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_smpl_frame_forward(void* stream_, const MossSmplFrame* f, float* A, float* rot_mats, float* joints,
	float* poseoff, float* shapeoff)
{
	MOSS_API_RANGE("moss_smpl_frame_forward");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!smpl_frame_valid(f) || !A || !f->J_template || (f->n_shapes > 0 && !f->J_shapedirs) || (poseoff && !f->posedirs) ||
		(shapeoff && f->n_shapes > 0 && !f->shapedirs)) return MOSS_ERR_INVALID_ARGUMENT;
	SmplFrameArgs a;
	a.V = f->V; a.n_shapes = f->n_shapes; a.shape_stride = f->shape_stride;
	a.shapedirs = f->shapedirs; a.posedirs = f->posedirs; a.J_template = f->J_template; a.J_shapedirs = f->J_shapedirs;
	a.parents = f->parents; a.poses = f->poses; a.shapes = f->shapes; a.correct_Rs = f->correct_Rs;
	a.A = A; a.rot_mats = rot_mats; a.joints = joints; a.poseoff = poseoff; a.shapeoff = shapeoff;
	a.shape_blocks = shapeoff ? smpl_shape_blocks(f->V) : 0;
	const int blocks = 1 + a.shape_blocks + (poseoff ? smpl_pose_blocks(f->V) : 0);
	{ MOSS_STAGE(MOSS_STAGE_SMPL_FRAME, stream);
		smpl_frame_forward_kernel<<<blocks, SMPL_THREADS, 0, stream>>>(a); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_smpl_frame_backward(void* stream_, const MossSmplFrame* f, const float* A, const float* joints,
	const float* d_A, const float* d_poseoff, float* d_poses, float* d_correct_Rs)
{
	MOSS_API_RANGE("moss_smpl_frame_backward");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!smpl_frame_valid(f) || !A || !joints || !d_poses || (d_poseoff && !f->posedirs) || !f->scratch ||
		f->scratch_bytes < moss_smpl_frame_scratch_bytes(f->V) || ((uintptr_t)f->scratch & 15)) return MOSS_ERR_INVALID_ARGUMENT;
	int dev = 0;                            // opt in to > 48 KB of dynamic shared memory (the attribute is per device)
	MOSS_CUDA_TRY(cudaGetDevice(&dev));
	if (dev >= 0 && dev < MAX_DEVICES && !g_smpl_attr_set[dev].load(std::memory_order_acquire)) {
		MOSS_CUDA_TRY(cudaFuncSetAttribute(smpl_frame_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMPL_BWD_DYN_SMEM));
		g_smpl_attr_set[dev].store(true, std::memory_order_release);
	}
	SmplFrameBwdArgs a;
	a.V = f->V; a.posedirs = f->posedirs; a.parents = f->parents; a.poses = f->poses; a.correct_Rs = f->correct_Rs;
	a.A = A; a.joints = joints; a.d_A = d_A; a.d_poseoff = d_poseoff; a.d_poses = d_poses; a.d_correct_Rs = d_correct_Rs;
	a.ticket = (unsigned*)f->scratch; a.chain_dR = (float*)((char*)f->scratch + 16); a.partial = a.chain_dR + NJ * 9;
	a.pose_blocks = d_poseoff ? smpl_bwd_blocks(f->V) : 0;
	{ MOSS_STAGE(MOSS_STAGE_SMPL_FRAME_BWD, stream);
		smpl_frame_backward_kernel<<<1 + a.pose_blocks, SMPL_THREADS, a.pose_blocks ? SMPL_BWD_DYN_SMEM : 0, stream>>>(a); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

// ---- image-space losses ------------------------------------------------------------------------------------------------
// blocks of the launch: the tile grid is aligned to the crop's corner (loss.cuh), hence one extra column and row
static long long loss_tiles(int W, int H) { return (long long)((W + LOSS_T - 1) / LOSS_T + 1) * (long long)((H + LOSS_T - 1) / LOSS_T + 1); }
// few tiles (<= two per SM): one block per (tile, colour channel), 16 warps each; many tiles: one 8-warp block per tile
static bool loss_channel_split(int W, int H) { return loss_tiles(W, H) <= LOSS_CSPLIT_MAX_TILES; }
static long long loss_blocks(int W, int H) { return loss_tiles(W, H) * (loss_channel_split(W, H) ? 3 : 1); }
static cudaError_t loss_set_attributes()
{
	// > 48 KB of dynamic shared memory is opt-in; two resident blocks need 174 KB, so ask for the largest carve-out
	cudaError_t e = cudaFuncSetAttribute(image_losses_kernel<256, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LOSS_DYN_SMEM);
	if (e == cudaSuccess) e = cudaFuncSetAttribute(image_losses_kernel<256, false>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
	if (e == cudaSuccess) e = cudaFuncSetAttribute(image_losses_kernel<LOSS_SPLIT_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LOSS_DYN_SMEM);
	if (e == cudaSuccess) e = cudaFuncSetAttribute(image_losses_kernel<LOSS_SPLIT_THREADS, true>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
	return e;
}
static void loss_launch(const LossArgs& a, cudaStream_t stream)
{
	const dim3 tiles((a.W + LOSS_T - 1) / LOSS_T + 1, (a.H + LOSS_T - 1) / LOSS_T + 1);
	if (loss_channel_split(a.W, a.H)) image_losses_kernel<LOSS_SPLIT_THREADS, true><<<dim3(tiles.x, tiles.y, 3), LOSS_SPLIT_THREADS, LOSS_DYN_SMEM, stream>>>(a);
	else image_losses_kernel<256, false><<<tiles, 256, LOSS_DYN_SMEM, stream>>>(a);
}
size_t moss_image_losses_scratch_bytes(int W, int H)
{
	if (W < 1 || H < 1) return 16;
	return 16 + (size_t)loss_blocks(W, H) * 4 * sizeof(float);
}
size_t moss_sizeof_image_loss(void) { return sizeof(MossImageLoss); }

// The 11-tap window as utils/loss_utils.py:47-49 builds it: float32 of exp(-(x-5)^2 / (2 * 1.5^2)), divided by their
// float32 sum.  S3IM taps of replica r on (x-1, x, x+1): the window entries that fall on each source pixel when every
// pixel is repeated ten times (stretched position 10x + r + d, d = -5..5).
static void loss_window(float* g, float (*s3)[3])
{
	float sum = 0.f;
	for (int x = 0; x < LOSS_WIN; x++) { g[x] = (float)exp(-(double)((x - LOSS_R) * (x - LOSS_R)) / (2.0 * 1.5 * 1.5)); sum += g[x]; }
	for (int x = 0; x < LOSS_WIN; x++) g[x] /= sum;
	for (int r = 0; r < LOSS_S3_REP; r++) {
		s3[r][0] = s3[r][1] = s3[r][2] = 0.f;
		for (int d = -LOSS_R; d <= LOSS_R; d++) {
			const int pos = r + d;
			s3[r][pos < 0 ? 0 : (pos < LOSS_S3_REP ? 1 : 2)] += g[d + LOSS_R];
		}
	}
}

int moss_image_losses(void* stream_, const MossImageLoss* k, float* losses, float* dL_dimage, float* dL_dalpha)
{
	MOSS_API_RANGE("moss_image_losses");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!k || k->W < 1 || k->H < 1 || !k->image || !k->gt_image || !k->rect || !losses || !k->scratch ||
		k->scratch_bytes < moss_image_losses_scratch_bytes(k->W, k->H) || ((uintptr_t)k->scratch & 15)) return MOSS_ERR_INVALID_ARGUMENT;
	if (k->w_mask != 0.f && (!k->alpha || !k->bkgd_mask)) return MOSS_ERR_INVALID_ARGUMENT;
	int dev = 0;                            // opt in to > 48 KB of dynamic shared memory (the attribute is per device)
	MOSS_CUDA_TRY(cudaGetDevice(&dev));
	if (dev >= 0 && dev < MAX_DEVICES && !g_loss_attr_set[dev].load(std::memory_order_acquire)) {
		MOSS_CUDA_TRY(loss_set_attributes());
		g_loss_attr_set[dev].store(true, std::memory_order_release);
	}
	LossArgs a;
	a.W = k->W; a.H = k->H; a.image = k->image; a.alpha = k->alpha; a.gt = k->gt_image; a.bkgd = k->bkgd_mask; a.bound = k->bound_mask;
	a.rect = k->rect; a.w_l1 = k->w_l1; a.w_mask = k->w_mask; a.w_ssim = k->w_ssim; a.w_s3im = k->w_s3im;
	loss_window(a.g, a.s3);
	a.dL_dimage = dL_dimage; a.dL_dalpha = dL_dalpha; a.losses = losses;
	a.ticket = (unsigned*)k->scratch; a.partial = (float*)((char*)k->scratch + 16);
	{ MOSS_STAGE(MOSS_STAGE_IMAGE_LOSSES, stream);
		loss_launch(a, stream); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

// ---- image losses without any allocation (no reference counterpart) ----
// The scratch of image_losses_kernel (ticket at +0, per-tile partial sums at +16) lives in a file-scope device array.
// In host code `g_loss_static_scratch` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device with that device current, and is what the kernel receives.
static __device__ __align__(256) unsigned char g_loss_static_scratch[MOSS_IMAGE_LOSSES_STATIC_SCRATCH_BYTES];
static std::atomic<void*> g_loss_static_ptr[MAX_DEVICES];
// ceil(W/32) * ceil(H/32) <= MAX_TILES bounds the launch's (ceil(W/32) + 1) * (ceil(H/32) + 1) blocks by 2 * MAX_TILES + 2
static_assert(MOSS_IMAGE_LOSSES_STATIC_SCRATCH_BYTES >= 16 + 16 * (2 * MOSS_IMAGE_LOSSES_STATIC_MAX_TILES + 2), "static scratch holds ticket + partial sums");
static_assert(3 * LOSS_CSPLIT_MAX_TILES <= MOSS_IMAGE_LOSSES_STATIC_MAX_TILES, "the per-channel blocks of a small image fit the static scratch too");

int moss_image_losses_static(void* stream_, const MossImageLoss* k, float* losses, float* dL_dimage, float* dL_dalpha)
{
	MOSS_API_RANGE("moss_image_losses_static");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!k || k->W < 1 || k->H < 1 || !k->image || !k->gt_image || !k->rect || !losses) return MOSS_ERR_INVALID_ARGUMENT;
	if (k->w_mask != 0.f && (!k->alpha || !k->bkgd_mask)) return MOSS_ERR_INVALID_ARGUMENT;
	const long long tiles = (long long)((k->W + LOSS_T - 1) / LOSS_T) * (long long)((k->H + LOSS_T - 1) / LOSS_T);
	if (tiles > MOSS_IMAGE_LOSSES_STATIC_MAX_TILES) return MOSS_ERR_STATIC_SCRATCH_LIMIT;
	int dev = 0;
	MOSS_CUDA_TRY(cudaGetDevice(&dev));
	void* scratch = (dev >= 0 && dev < MAX_DEVICES) ? g_loss_static_ptr[dev].load(std::memory_order_acquire) : nullptr;
	if (!scratch) {                         // set-up, once per device: the array's address ON THE CURRENT DEVICE
		MOSS_CUDA_TRY(cudaGetSymbolAddress((void**)&scratch, g_loss_static_scratch));
		if (!scratch) return MOSS_ERR_CUDA;
		if (dev >= 0 && dev < MAX_DEVICES) g_loss_static_ptr[dev].store(scratch, std::memory_order_release);
	}
	if (dev >= 0 && dev < MAX_DEVICES && !g_loss_attr_set[dev].load(std::memory_order_acquire)) {
		MOSS_CUDA_TRY(loss_set_attributes());
		g_loss_attr_set[dev].store(true, std::memory_order_release);
	}
	LossArgs a;
	a.W = k->W; a.H = k->H; a.image = k->image; a.alpha = k->alpha; a.gt = k->gt_image; a.bkgd = k->bkgd_mask; a.bound = k->bound_mask;
	a.rect = k->rect; a.w_l1 = k->w_l1; a.w_mask = k->w_mask; a.w_ssim = k->w_ssim; a.w_s3im = k->w_s3im;
	loss_window(a.g, a.s3);
	a.dL_dimage = dL_dimage; a.dL_dalpha = dL_dalpha; a.losses = losses;
	a.ticket = (unsigned*)scratch; a.partial = (float*)((char*)scratch + 16);
	// the array keeps the last call's contents: the ticket must be zero on entry (the partial sums are fully rewritten)
	MOSS_CUDA_TRY(cudaMemsetAsync(scratch, 0, 16, stream));
	g_launches.fetch_add(1, std::memory_order_relaxed);          // no MOSS_STAGE here: the profiler creates events
	loss_launch(a, stream);
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_mask_bounds(void* stream_, int W, int H, const float* mask, int* out)
{
	MOSS_API_RANGE("moss_mask_bounds");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (W < 1 || H < 1 || !mask || !out) return MOSS_ERR_INVALID_ARGUMENT;
	MOSS_CUDA_TRY(cudaMemsetAsync(out, 0x7f, 2 * sizeof(int), stream));
	MOSS_CUDA_TRY(cudaMemsetAsync(out + 2, 0x80, 2 * sizeof(int), stream));
	MOSS_CUDA_TRY(cudaMemsetAsync(out + 4, 0, 4 * sizeof(int), stream));
	const size_t N = (size_t)W * H;
	const int blocks = (int)std::min<size_t>((N + 255) / 256, 148 * 8);
	{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		mask_bounds_kernel<<<blocks, 256, 0, stream>>>(W, H, mask, out); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_scale_gradients(void* stream_, const float* scale_dev, float* a, size_t na, float* b, size_t nb)
{
	cudaStream_t stream = (cudaStream_t)stream_;
	if (!scale_dev || (na && !a) || (nb && !b)) return MOSS_ERR_INVALID_ARGUMENT;
	if (na + nb == 0) return MOSS_OK;
	const int blocks = (int)std::min<size_t>((na + nb + 255) / 256, 148 * 8);
	{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		scale_inplace_kernel<<<blocks, 256, 0, stream>>>(scale_dev, a, na, b, nb); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_nearest_vertex(void* stream_, int P, const float* xyz, int V, const float* verts, int* vert_ids, float* dist)
{
	MOSS_API_RANGE("moss_nearest_vertex");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (P < 0 || V < 1) return MOSS_ERR_INVALID_ARGUMENT;
	if (P == 0) return MOSS_OK;
	if (!xyz || !verts || !vert_ids) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_NEAREST, stream);
		nearest_vertex_kernel<<<(P + 31) / 32, NEAR_THREADS, 0, stream>>>(P, xyz, V, verts, vert_ids, dist); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

size_t moss_vertex_grid_bytes(int V) { return V < 1 ? 0 : vertex_grid_bytes(V); }

int moss_vertex_grid_build(void* stream_, int V, const float* verts, void* grid, size_t grid_bytes)
{
	MOSS_API_RANGE("moss_vertex_grid_build");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (V < 1 || !verts || !grid || ((uintptr_t)grid & 15) || grid_bytes < vertex_grid_bytes(V)) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_OTHER, stream);
		vertex_grid_build_kernel<<<1, GRID_BUILD_THREADS, 0, stream>>>(V, verts, vertex_grid_view(grid, V)); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_nearest_vertex_grid(void* stream_, int P, const float* xyz, int V, const void* grid, size_t grid_bytes,
	int* vert_ids, float* dist)
{
	MOSS_API_RANGE("moss_nearest_vertex_grid");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (P < 0 || V < 1 || !grid || ((uintptr_t)grid & 15) || grid_bytes < vertex_grid_bytes(V)) return MOSS_ERR_INVALID_ARGUMENT;
	if (P == 0) return MOSS_OK;
	if (!xyz || !vert_ids) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_NEAREST, stream);
		nearest_vertex_grid_kernel<<<(unsigned)(((size_t)P * GRID_GW + GRID_QUERY_THREADS - 1) / GRID_QUERY_THREADS), GRID_QUERY_THREADS, 0, stream>>>(P, xyz, vertex_grid_view(const_cast<void*>(grid), V), vert_ids, dist); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

int moss_densify_stats(void* stream_, int P, const int* radii, const float* dL_dmean2D, float* xyz_gradient_accum,
	float* denom, float* max_radii2D)
{
	MOSS_API_RANGE("moss_densify_stats");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (P < 0) return MOSS_ERR_INVALID_ARGUMENT;
	if (P == 0) return MOSS_OK;
	if (!radii || (xyz_gradient_accum && !dL_dmean2D)) return MOSS_ERR_INVALID_ARGUMENT;
	{ MOSS_STAGE(MOSS_STAGE_DENSIFY_STATS, stream);
		densify_stats_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, radii, dL_dmean2D, xyz_gradient_accum, denom, max_radii2D); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

size_t moss_sizeof_adam_group(void) { return sizeof(MossAdamGroup); }

int moss_adamw_step(void* stream_, int n_groups, const MossAdamGroup* groups, double beta1, double beta2, double eps,
	double weight_decay)
{
	MOSS_API_RANGE("moss_adamw_step");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (n_groups < 0 || n_groups > ADAM_MAX_GROUPS || (n_groups && !groups)) return MOSS_ERR_INVALID_ARGUMENT;
	AdamGroups a;
	memset(&a, 0, sizeof(a));
	unsigned long long chunks = 0;
	for (int i = 0; i < n_groups; i++) {
		const MossAdamGroup& g = groups[i];
		if (g.numel == 0) continue;
		if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq || g.step < 1) return MOSS_ERR_INVALID_ARGUMENT;
		const int k = a.n++;
		a.param[k] = g.param; a.grad[k] = g.grad; a.exp_avg[k] = g.exp_avg; a.exp_avg_sq[k] = g.exp_avg_sq;
		a.numel[k] = g.numel;
		chunks += (g.numel + ADAM_CHUNK - 1) / ADAM_CHUNK;
		if (chunks > 0x7fffffffull) return MOSS_ERR_INVALID_ARGUMENT;
		a.chunk_end[k] = (unsigned)chunks;
		// the scalars in double, like the Python floats torch derives them from
		const double lr = g.lr;
		a.decay[k] = (float)(1.0 - lr * weight_decay);
		a.step_size[k] = (float)(lr / (1.0 - pow(beta1, (double)g.step)));
		a.bc2_sqrt[k] = (float)sqrt(1.0 - pow(beta2, (double)g.step));
	}
	if (a.n == 0) return MOSS_OK;
	a.w1 = (float)(1.0 - beta1); a.beta2 = (float)beta2; a.w2 = (float)(1.0 - beta2); a.eps = (float)eps;
	{ MOSS_STAGE(MOSS_STAGE_ADAMW, stream);
		adamw_kernel<<<(unsigned)chunks, 256, 0, stream>>>(a); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

// ---- packed-gradient all-reduce over mapped peer memory -------------------------------------------------------
size_t moss_peer_signal_bytes(int world) { return sizeof(uint32_t) * 2 * 8 * (size_t)(world > 0 ? world : 1); }   // minimum: 8 blocks

int moss_peer_allreduce(void* stream_, int rank, int world, void* const* buffer_ptrs_dev, void* const* signal_ptrs_dev,
	size_t signal_pad_bytes, size_t offset_floats, size_t numel, unsigned epoch, unsigned* error_word_dev,
	void* multicast_ptr)
{
	MOSS_API_RANGE("moss_peer_allreduce");
	cudaStream_t stream = (cudaStream_t)stream_;
	if (world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || !buffer_ptrs_dev || !signal_ptrs_dev ||
		!error_word_dev || (offset_floats & 3) || signal_pad_bytes < moss_peer_signal_bytes(world))
		return MOSS_ERR_INVALID_ARGUMENT;
	if (numel == 0) return MOSS_OK;
	// every rank derives the same grid from (numel, world, pad size): the flag slots are indexed by it
	const size_t per = (numel / 4 + world - 1) / world;
	size_t blocks = (per + (size_t)PEER_THREADS * PEER_ILP - 1) / ((size_t)PEER_THREADS * PEER_ILP);
	const size_t pad_blocks = signal_pad_bytes / (sizeof(uint32_t) * 2 * (size_t)world);
	if (blocks > pad_blocks) blocks = pad_blocks;
	if (blocks > PEER_MAX_BLOCKS) blocks = PEER_MAX_BLOCKS;
	if (blocks < 1) blocks = 1;
	{ MOSS_STAGE(MOSS_STAGE_PEER_REDUCE, stream);
	peer_allreduce_kernel<<<(unsigned)blocks, PEER_THREADS, 0, stream>>>((float* const*)buffer_ptrs_dev, (uint32_t* const*)signal_ptrs_dev,
		rank, world, offset_floats, numel, epoch, error_word_dev, (float*)multicast_ptr); }
	MOSS_CHECK_LAUNCH(stream, 0);
	return MOSS_OK;
}

#ifdef MOSS_TILE_TIMING
int moss_debug_tile_timing(long long* dst, int tiles)
{
	cudaDeviceSynchronize();
	return (int)cudaMemcpyFromSymbol(dst, moss::g_tile_timing, sizeof(long long) * 8 * (size_t)tiles);
}
int moss_debug_tile_detail(long long* dst, int tiles)
{
	cudaDeviceSynchronize();
	return (int)cudaMemcpyFromSymbol(dst, moss::g_tile_detail, sizeof(long long) * 16 * (size_t)tiles);
}
#endif

} // extern "C"
