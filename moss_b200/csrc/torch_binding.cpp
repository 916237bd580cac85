// Compiled autograd binding of the drop-in API over the C ABI (include/moss_splat.h).
//
// The reference's host side is a pybind module + torch glue in C++ (DGR/ext.cpp:15-19, DGR/rasterize_points.cu:35-206)
// under a Python autograd.Function (DGR/diff_gaussian_rasterization/__init__.py:44-158).  At ZJU size one fwd+bwd is
// ~0.16 ms of kernels, so the host path IS the critical path of the public API: this file is the same glue written
// once in C++ -- argument checks, workspace allocation through torch's caching allocator, the C-ABI calls and the
// autograd node -- so that `GaussianRasterizer(...)` costs a few tens of microseconds of host time instead of ~150.
// Nothing here computes: every kernel lives in libmoss_splat.so; torch provides memory, the stream and autograd.
#include <torch/extension.h>
#include <c10/cuda/CUDAStream.h>
#include <c10/cuda/CUDAGuard.h>
#include <mutex>
#include <unordered_map>
#include <vector>
#include "../../include/moss_splat.h"

namespace {

using torch::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

inline const float* fptr(const Tensor& t) { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; }

Tensor f32c(const Tensor& t, const char* name)
{
	if (!t.defined() || t.numel() == 0) return t;
	TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor");
	TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, " must be float32");
	return t.contiguous();
}

inline int64_t align256(int64_t n) { return (n + 255) / 256 * 256; }

// remembered instance counts so that the binning workspace is right-sized without a host sync
std::mutex g_hint_mu;
std::unordered_map<uint64_t, int64_t> g_hint;
uint64_t hint_key(int64_t P, int64_t W, int64_t H, int64_t V, int dev)
{
	return ((uint64_t)P * 1000003ull + (uint64_t)W) * 1000003ull + (uint64_t)H * 4099ull + (uint64_t)V * 67ull + (uint64_t)dev;
}

// data-parallel gradient sink (moss_b200.set_gradient_sink): the backward writes into this symmetric buffer with the
// exchange fused in (include/moss_splat.h: moss_rasterize_backward_push)
struct Sink {
	Tensor tensor;                       // this rank's packed gradient buffer (symmetric memory)
	std::vector<void*> grad_ptrs, stage_ptrs, pad_ptrs;
	MossPush push;
	bool on = false;
};
Sink g_sink;

void check(int status, const char* what)
{
	if (status == MOSS_OK) return;
	std::string msg = std::string(what) + ": " + moss_error_string(status);
	if (status == MOSS_ERR_CUDA) msg += " (cudaError " + std::to_string(moss_last_cuda_error()) + ")";
	TORCH_CHECK(false, msg);
}

struct Views {
	std::vector<MossView> v;
	std::vector<Tensor> keep;
};

// settings of V views: per-view tensors bg, viewmatrix, projmatrix, campos and floats tanfovx, tanfovy
Views make_views(const std::vector<Tensor>& bg, const std::vector<Tensor>& view, const std::vector<Tensor>& proj,
	const std::vector<Tensor>& campos, const std::vector<double>& tanfovx, const std::vector<double>& tanfovy)
{
	Views o;
	const size_t V = view.size();
	TORCH_CHECK(V >= 1 && V <= MOSS_MAX_VIEWS, "a batched call takes 1..", MOSS_MAX_VIEWS, " views, got ", V);
	TORCH_CHECK(bg.size() == V && proj.size() == V && campos.size() == V && tanfovx.size() == V && tanfovy.size() == V,
		"per-view settings must have one entry per view");
	o.v.resize(V);
	for (size_t i = 0; i < V; i++) {
		Tensor b = f32c(bg[i], "bg"), vm = f32c(view[i], "viewmatrix"), pm = f32c(proj[i], "projmatrix"), cp = f32c(campos[i], "campos");
		o.v[i].bg = fptr(b); o.v[i].viewmatrix = fptr(vm); o.v[i].projmatrix = fptr(pm); o.v[i].campos = fptr(cp);
		o.v[i].tanfovx = (float)tanfovx[i]; o.v[i].tanfovy = (float)tanfovy[i];
		o.keep.push_back(b); o.keep.push_back(vm); o.keep.push_back(pm); o.keep.push_back(cp);
	}
	return o;
}

struct GradLayout { int64_t means3D, cov3D, sh, scales, rot, opacity, means2D, colors, conic, param_numel, total; };
// the packed gradient buffer of moss_b200/rasterizer.py::grad_layout (every sub-array 16 B-aligned)
GradLayout grad_layout(int64_t P, int64_t M, bool has_sh, bool has_sr, int64_t V)
{
	auto pad = [](int64_t n) { return (n + 3) / 4 * 4; };
	GradLayout l; int64_t o = 0;
	l.means3D = o; o += pad(3 * P);
	l.cov3D = o; o += pad(6 * P);
	l.sh = o; o += pad(has_sh ? 3 * M * P : 0);
	l.scales = o; o += pad(has_sr ? 3 * P : 0);
	l.rot = o; o += pad(has_sr ? 4 * P : 0);
	l.opacity = o; o += pad(P);
	if (V > 1) l.param_numel = o;
	l.means2D = o; o += pad(3 * P * V);
	if (V == 1) l.param_numel = o;
	l.colors = o; o += pad(3 * P);
	l.conic = o; o += pad(4 * P * V);
	l.total = o;
	return l;
}

// everything that is not a differentiable input travels as ONE opaque argument of the autograd node
struct Cfg {
	std::vector<Tensor> bg, viewmatrix, projmatrix, campos;
	std::vector<double> tanfovx, tanfovy;
	int64_t H, W, sh_degree, flags;
	double scale_modifier;
	bool prefiltered, debug, batched, want_ckpt;
};

class RasterizeFn : public torch::autograd::Function<RasterizeFn> {
public:
	static variable_list forward(AutogradContext* ctx, Tensor means3D, Tensor means2D, Tensor sh, Tensor colors_precomp,
		Tensor opacities, Tensor scales, Tensor rotations, Tensor cov3D_precomp, const Cfg& cfg)
	{
		(void)means2D;
		const auto& bg = cfg.bg; const auto& viewmatrix = cfg.viewmatrix; const auto& projmatrix = cfg.projmatrix; const auto& campos = cfg.campos;
		const auto& tanfovx = cfg.tanfovx; const auto& tanfovy = cfg.tanfovy;
		const int64_t H = cfg.H, W = cfg.W, sh_degree = cfg.sh_degree;
		int64_t flags = cfg.flags;
		const double scale_modifier = cfg.scale_modifier;
		const bool prefiltered = cfg.prefiltered, debug = cfg.debug, batched = cfg.batched, want_ckpt = cfg.want_ckpt;
		TORCH_CHECK(means3D.dim() == 2 && means3D.size(1) == 3, "means3D must have dimensions (num_points, 3)");   // DGR/rasterize_points.cu:57-59
		TORCH_CHECK(means3D.is_cuda(), "moss_b200 rasterizer needs CUDA tensors (there is no CPU path)");
		const c10::cuda::CUDAGuard guard(means3D.device());
		means3D = f32c(means3D, "means3D"); sh = f32c(sh, "sh"); colors_precomp = f32c(colors_precomp, "colors_precomp");
		opacities = f32c(opacities, "opacities"); scales = f32c(scales, "scales"); rotations = f32c(rotations, "rotations");
		cov3D_precomp = f32c(cov3D_precomp, "cov3Ds_precomp");
		Views views = make_views(bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy);
		const int64_t V = (int64_t)views.v.size();
		TORCH_CHECK(batched || V == 1, "a single-view call takes one view");
		const int64_t P = means3D.size(0);
		const int64_t M = sh.numel() ? sh.size(1) : 0;
		if (debug) flags |= MOSS_FLAG_DEBUG;
		const int export_keys = (flags & MOSS_FLAG_EXPORT_KEYS) ? 1 : 0;
		const auto dev = means3D.device();
		const auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
		const auto u8 = torch::TensorOptions().dtype(torch::kUInt8).device(dev);

		// three independent output tensors, like the reference (DGR/rasterize_points.cu:68-71)
		std::vector<int64_t> lead;
		if (batched) lead.push_back(V);
		auto shape = [&](std::initializer_list<int64_t> tail) { std::vector<int64_t> s = lead; s.insert(s.end(), tail); return s; };
		Tensor color = torch::empty(shape({3, H, W}), f32);
		Tensor depth = torch::empty(shape({1, H, W}), f32);
		Tensor alpha = torch::empty(shape({1, H, W}), f32);
		Tensor radii = torch::empty(shape({P}), f32.dtype(torch::kInt32));
		const int64_t gbytes = align256((int64_t)moss_geom_workspace_bytes_batch((int)P, (int)V));
		const int64_t ibytes = (int64_t)moss_image_workspace_bytes_batch((int)W, (int)H, (int)V);
		Tensor ws = torch::empty({gbytes + ibytes}, u8);

		const uint64_t key = hint_key(P, W, H, V, dev.index());
		int64_t cap = 0;
		{ std::lock_guard<std::mutex> lk(g_hint_mu); auto it = g_hint.find(key); if (it != g_hint.end()) cap = it->second; }
		cap = std::max<int64_t>((int64_t)(cap * 1.25) + 4096, (4 * P + 4096) * V);

		MossScene sc;
		memset(&sc, 0, sizeof(sc));
		sc.P = (int)P; sc.D = (int)sh_degree; sc.M = (int)M; sc.W = (int)W; sc.H = (int)H;
		sc.means3D = fptr(means3D); sc.shs = fptr(sh); sc.colors_precomp = fptr(colors_precomp); sc.opacities = fptr(opacities);
		sc.scales = fptr(scales); sc.rotations = fptr(rotations); sc.cov3D_precomp = fptr(cov3D_precomp);
		sc.scale_modifier = (float)scale_modifier; sc.prefiltered = prefiltered ? 1 : 0;
		if (!batched) {
			sc.bg = views.v[0].bg; sc.viewmatrix = views.v[0].viewmatrix; sc.projmatrix = views.v[0].projmatrix; sc.campos = views.v[0].campos;
			sc.tanfovx = views.v[0].tanfovx; sc.tanfovy = views.v[0].tanfovy;
		}
		sc.geom_ws = ws.data_ptr(); sc.geom_bytes = (size_t)gbytes;
		sc.image_ws = (char*)ws.data_ptr() + gbytes; sc.image_bytes = (size_t)ibytes;
		void* stream = (void*)c10::cuda::getCurrentCUDAStream(dev.index()).stream();
		Tensor binning, ckpt;
		int nr = 0;
		unsigned call_flags = (unsigned)flags;
		for (;;) {
			const int64_t bbytes = (int64_t)moss_binning_workspace_bytes_ex(cap, export_keys);
			binning = torch::empty({bbytes}, u8);
			sc.binning_ws = binning.data_ptr(); sc.binning_bytes = (size_t)bbytes;
			if (want_ckpt) {
				const int64_t cbytes = (int64_t)moss_ckpt_workspace_bytes_batch(moss_binning_capacity((size_t)bbytes, export_keys), (int)W, (int)H, (int)V);
				ckpt = torch::empty({cbytes}, u8);
				sc.ckpt_ws = ckpt.data_ptr(); sc.ckpt_bytes = (size_t)cbytes;
			}
			int status;
			if (batched)
				status = moss_rasterize_forward_batch(stream, &sc, (int)V, views.v.data(), call_flags, color.data_ptr<float>(),
					depth.data_ptr<float>(), alpha.data_ptr<float>(), P ? radii.data_ptr<int>() : nullptr, &nr);
			else
				status = moss_rasterize_forward(stream, &sc, call_flags, color.data_ptr<float>(), depth.data_ptr<float>(),
					alpha.data_ptr<float>(), P ? radii.data_ptr<int>() : nullptr, &nr);
			if (status == MOSS_ERR_BINNING_OVERFLOW) {
				cap = (int64_t)(nr * 1.25) + 4096;
				call_flags = (unsigned)flags | MOSS_FLAG_REBIN;
				continue;
			}
			check(status, "moss_rasterize_forward");
			break;
		}
		if (nr >= 0) { std::lock_guard<std::mutex> lk(g_hint_mu); g_hint[key] = nr; }

		// The workspaces and the inputs whose device pointers the backward needs go through save_for_backward, like the
		// reference's three state tensors (DGR/diff_gaussian_rasterization/__init__.py:97): autograd frees them right
		// after the backward unless retain_graph is set.
		variable_list saved = {radii, ws, binning, ckpt.defined() ? ckpt : Tensor(), means3D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp};
		for (auto& t : views.keep) saved.push_back(t);
		ctx->save_for_backward(saved);
		ctx->saved_data["V"] = V; ctx->saved_data["H"] = H; ctx->saved_data["W"] = W; ctx->saved_data["D"] = sh_degree;
		ctx->saved_data["M"] = M; ctx->saved_data["P"] = P;
		ctx->saved_data["scale_modifier"] = scale_modifier; ctx->saved_data["prefiltered"] = prefiltered;
		ctx->saved_data["batched"] = batched; ctx->saved_data["flags"] = flags; ctx->saved_data["num_rendered"] = (int64_t)nr;
		ctx->saved_data["tanfovx"] = tanfovx; ctx->saved_data["tanfovy"] = tanfovy;
		ctx->mark_non_differentiable({radii});
		ctx->set_materialize_grads(false);      // an output nobody differentiates arrives undefined, not as a zero image
		return {color, radii, depth, alpha};
	}

	static variable_list backward(AutogradContext* ctx, variable_list grad_out)
	{
		const variable_list saved = ctx->get_saved_variables();      // raises the usual error when the graph was freed
		const Tensor& radii = saved[0]; const Tensor& ws = saved[1]; const Tensor& binning = saved[2]; const Tensor& ckpt = saved[3];
		const Tensor& means3D = saved[4]; const Tensor& sh = saved[5]; const Tensor& colors_precomp = saved[6];
		const Tensor& opacities = saved[7]; const Tensor& scales = saved[8]; const Tensor& rotations = saved[9];
		const Tensor& cov3D_precomp = saved[10];
		const int64_t V = ctx->saved_data["V"].toInt(), H = ctx->saved_data["H"].toInt(), W = ctx->saved_data["W"].toInt();
		const int64_t P = ctx->saved_data["P"].toInt(), M = ctx->saved_data["M"].toInt();
		const bool batched = ctx->saved_data["batched"].toBool();
		const int64_t flags = ctx->saved_data["flags"].toInt();
		const auto tanfovx = ctx->saved_data["tanfovx"].toDoubleVector(), tanfovy = ctx->saved_data["tanfovy"].toDoubleVector();
		const c10::cuda::CUDAGuard guard(means3D.device());
		const auto dev = means3D.device();
		const auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(dev);

		std::vector<MossView> views((size_t)V);
		for (int64_t i = 0; i < V; i++) {
			views[i].bg = fptr(saved[11 + 4 * i]); views[i].viewmatrix = fptr(saved[12 + 4 * i]);
			views[i].projmatrix = fptr(saved[13 + 4 * i]); views[i].campos = fptr(saved[14 + 4 * i]);
			views[i].tanfovx = (float)tanfovx[i]; views[i].tanfovy = (float)tanfovy[i];
		}
		MossScene sc;
		memset(&sc, 0, sizeof(sc));
		sc.P = (int)P; sc.D = (int)ctx->saved_data["D"].toInt(); sc.M = (int)M; sc.W = (int)W; sc.H = (int)H;
		sc.means3D = fptr(means3D); sc.shs = fptr(sh); sc.colors_precomp = fptr(colors_precomp); sc.opacities = fptr(opacities);
		sc.scales = fptr(scales); sc.rotations = fptr(rotations); sc.cov3D_precomp = fptr(cov3D_precomp);
		sc.scale_modifier = (float)ctx->saved_data["scale_modifier"].toDouble(); sc.prefiltered = ctx->saved_data["prefiltered"].toBool() ? 1 : 0;
		if (!batched) {
			sc.bg = views[0].bg; sc.viewmatrix = views[0].viewmatrix; sc.projmatrix = views[0].projmatrix; sc.campos = views[0].campos;
			sc.tanfovx = views[0].tanfovx; sc.tanfovy = views[0].tanfovy;
		}
		const int64_t gbytes = align256((int64_t)moss_geom_workspace_bytes_batch((int)P, (int)V));
		sc.geom_ws = ws.data_ptr(); sc.geom_bytes = (size_t)gbytes;
		sc.image_ws = (char*)ws.data_ptr() + gbytes; sc.image_bytes = (size_t)(ws.numel() - gbytes);
		sc.binning_ws = binning.data_ptr(); sc.binning_bytes = (size_t)binning.numel();
		if (ckpt.defined() && ckpt.numel()) { sc.ckpt_ws = ckpt.data_ptr(); sc.ckpt_bytes = (size_t)ckpt.numel(); }

		std::vector<int64_t> lead;
		if (batched) lead.push_back(V);
		auto shape = [&](std::initializer_list<int64_t> tail) { std::vector<int64_t> s = lead; s.insert(s.end(), tail); return s; };
		auto up = [&](const Tensor& g, std::initializer_list<int64_t> tail, bool required) -> Tensor {
			if (!g.defined()) return required ? torch::zeros(shape(tail), f32) : Tensor();     // NULL: no gradient reaches this output
			Tensor t = f32c(g, "grad");
			const auto s = shape(tail);
			if (t.sizes() != c10::IntArrayRef(s)) t = t.expand(s).contiguous();
			return t;
		};
		const Tensor gc = up(grad_out[0], {3, H, W}, true), gd = up(grad_out[2], {1, H, W}, false), ga = up(grad_out[3], {1, H, W}, false);

		const bool has_sh = sh.numel() != 0, has_sr = scales.numel() != 0;
		const GradLayout l = grad_layout(P, M, has_sh, has_sr, V);
		const MossPush* push = nullptr;
		MossPush push_desc;
		Tensor buf;
		if (g_sink.on) {
			TORCH_CHECK(g_sink.tensor.numel() >= l.total, "gradient sink holds ", g_sink.tensor.numel(), " floats, this call needs ", l.total);
			TORCH_CHECK(g_sink.tensor.device() == dev, "gradient sink lives on another device");
			push_desc = g_sink.push;
			push_desc.numel = (size_t)l.param_numel;
			push = &push_desc;
			buf = g_sink.tensor;
		} else {
			// every row of every output is written by the backward preprocess kernel: nothing is zero-filled here
			buf = torch::empty({l.total}, f32);
		}
		auto view_of = [&](int64_t off, int64_t n, std::vector<int64_t> s) { return buf.narrow(0, off, n).view(s); };
		Tensor dmeans3D = view_of(l.means3D, 3 * P, {P, 3}), dcov3D = view_of(l.cov3D, 6 * P, {P, 6});
		Tensor dopacity = view_of(l.opacity, P, {P, 1}), dcolors = view_of(l.colors, 3 * P, {P, 3});
		Tensor dmeans2D = view_of(l.means2D, 3 * P * V, shape({P, 3})), dconic = view_of(l.conic, 4 * P * V, shape({P, 2, 2}));
		Tensor dsh = has_sh ? view_of(l.sh, 3 * M * P, {P, M, 3}) : torch::zeros({P, 0, 3}, f32);
		// absent inputs get an undefined gradient (the reference zero-fills two tensors per backward for them,
		// DGR/rasterize_points.cu:164-165: two fill launches nobody reads)
		Tensor dscales = has_sr ? view_of(l.scales, 3 * P, {P, 3}) : Tensor();
		Tensor drot = has_sr ? view_of(l.rot, 4 * P, {P, 4}) : Tensor();
		void* stream = (void*)c10::cuda::getCurrentCUDAStream(dev.index()).stream();
		Tensor tmp;
		size_t tbytes = 0;
		if (batched && V > 1) {
			tbytes = moss_batch_grad_workspace_bytes((int)P, (int)M, (int)V);
			tmp = torch::empty({(int64_t)tbytes}, f32.dtype(torch::kUInt8));
		}
		auto p = [](Tensor& t) -> float* { return t.numel() ? t.data_ptr<float>() : nullptr; };
		int status;
		if (P == 0) status = MOSS_OK;
		else if (push || batched)
			status = push ? moss_rasterize_backward_push(stream, &sc, (int)V, batched ? views.data() : nullptr, (unsigned)flags, radii.data_ptr<int>(),
					fptr(gc), fptr(gd), fptr(ga), p(dmeans2D), p(dconic), p(dopacity), p(dcolors), p(dmeans3D), p(dcov3D),
					has_sh ? p(dsh) : nullptr, has_sr ? p(dscales) : nullptr, has_sr ? p(drot) : nullptr, tmp.defined() ? tmp.data_ptr() : nullptr, tbytes, push)
				: moss_rasterize_backward_batch(stream, &sc, (int)V, views.data(), (unsigned)flags, radii.data_ptr<int>(),
					fptr(gc), fptr(gd), fptr(ga), p(dmeans2D), p(dconic), p(dopacity), p(dcolors), p(dmeans3D), p(dcov3D),
					has_sh ? p(dsh) : nullptr, has_sr ? p(dscales) : nullptr, has_sr ? p(drot) : nullptr, tmp.defined() ? tmp.data_ptr() : nullptr, tbytes);
		else
			status = moss_rasterize_backward(stream, &sc, (unsigned)flags, radii.data_ptr<int>(), fptr(gc), fptr(gd), fptr(ga),
				p(dmeans2D), p(dconic), p(dopacity), p(dcolors), p(dmeans3D), p(dcov3D),
				has_sh ? p(dsh) : nullptr, has_sr ? p(dscales) : nullptr, has_sr ? p(drot) : nullptr);
		check(status, "moss_rasterize_backward");
		// same order as DGR/diff_gaussian_rasterization/__init__.py:146-156, then one undefined entry per setting
		variable_list out = {dmeans3D, dmeans2D, dsh, dcolors, dopacity, dscales, drot, dcov3D, Tensor()};
		return out;
	}
};

std::vector<Tensor> rasterize(Tensor means3D, Tensor means2D, Tensor sh, Tensor colors_precomp, Tensor opacities, Tensor scales,
	Tensor rotations, Tensor cov3D_precomp, std::vector<Tensor> bg, std::vector<Tensor> viewmatrix, std::vector<Tensor> projmatrix,
	std::vector<Tensor> campos, std::vector<double> tanfovx, std::vector<double> tanfovy, int64_t H, int64_t W, double scale_modifier,
	int64_t sh_degree, bool prefiltered, bool debug, bool batched, bool want_ckpt, int64_t flags)
{
	Cfg cfg;
	cfg.bg = std::move(bg); cfg.viewmatrix = std::move(viewmatrix); cfg.projmatrix = std::move(projmatrix); cfg.campos = std::move(campos);
	cfg.tanfovx = std::move(tanfovx); cfg.tanfovy = std::move(tanfovy);
	cfg.H = H; cfg.W = W; cfg.sh_degree = sh_degree; cfg.flags = flags; cfg.scale_modifier = scale_modifier;
	cfg.prefiltered = prefiltered; cfg.debug = debug; cfg.batched = batched; cfg.want_ckpt = want_ckpt;
	return RasterizeFn::apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp, cfg);
}

// tensor: the symmetric packed gradient buffer; the pointer lists are every rank's buffer / staging area / signal pad as
// mapped in this process (parallel.PeerGradBuffer); pass an undefined tensor to detach the sink
void set_gradient_sink(c10::optional<Tensor> tensor, int64_t rank, int64_t world, std::vector<int64_t> grad_ptrs,
	std::vector<int64_t> stage_ptrs, std::vector<int64_t> pad_ptrs, int64_t pad_bytes, int64_t multicast_ptr, int64_t error_ptr)
{
	g_sink.on = false;
	g_sink.tensor = Tensor();
	if (!tensor.has_value() || !tensor->defined()) return;
	TORCH_CHECK((int64_t)grad_ptrs.size() == world && (int64_t)stage_ptrs.size() == world && (int64_t)pad_ptrs.size() == world,
		"one pointer per rank expected");
	g_sink.tensor = *tensor;
	g_sink.grad_ptrs.clear(); g_sink.stage_ptrs.clear(); g_sink.pad_ptrs.clear();
	for (int64_t i = 0; i < world; i++) {
		g_sink.grad_ptrs.push_back((void*)grad_ptrs[i]); g_sink.stage_ptrs.push_back((void*)stage_ptrs[i]); g_sink.pad_ptrs.push_back((void*)pad_ptrs[i]);
	}
	memset(&g_sink.push, 0, sizeof(MossPush));
	g_sink.push.rank = (int)rank; g_sink.push.world = (int)world;
	g_sink.push.grad_base = tensor->data_ptr<float>();
	g_sink.push.grad_ptrs = g_sink.grad_ptrs.data(); g_sink.push.stage_ptrs = g_sink.stage_ptrs.data(); g_sink.push.signal_ptrs = g_sink.pad_ptrs.data();
	g_sink.push.signal_pad_bytes = (size_t)pad_bytes;
	g_sink.push.multicast_grad = (void*)multicast_ptr;
	g_sink.push.error_word_dev = (unsigned*)error_ptr;
	g_sink.on = true;
}

} // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m)
{
	m.doc() = "compiled autograd binding of moss_b200's drop-in rasterizer API over the C ABI (include/moss_splat.h)";
	m.def("rasterize", &rasterize, "forward of the autograd node; returns [color, radii, depth, alpha]");
	m.def("set_gradient_sink", &set_gradient_sink);
	m.def("abi_version", []() { return moss_abi_version(); });
}
