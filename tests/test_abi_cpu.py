"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/moss_splat.h declares.
No compute is launched (there is no GPU here); only host-side helpers are exercised."""
import ctypes as C
import os
import re

import pytest
import torch

from moss_b200 import _abi, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _abi.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "moss_splat.h")).read()
    declared = set(re.findall(r"\b(moss_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_abi.SYMBOLS), declared ^ set(_abi.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s


def test_versions_and_strings(lib):
    assert lib.moss_abi_version() == 10
    assert lib.moss_error_string(0) == b"ok"
    assert b"binning" in lib.moss_error_string(_abi.MOSS_ERR_BINNING_OVERFLOW)


def test_workspace_sizes_monotone(lib):
    prev = 0
    for P in (0, 1, 1000, 20000, 1000000):
        b = lib.moss_geom_workspace_bytes(P)
        assert b >= prev and b >= 70 * P
        prev = b
    assert lib.moss_image_workspace_bytes(512, 512) >= 8 * 512 * 512
    for cap in (0, 1, 4096, 123457, 10_000_000):
        for ex in (0, 1):
            nb = lib.moss_binning_workspace_bytes_ex(cap, ex)
            assert lib.moss_binning_capacity(nb, ex) >= cap
            assert lib.moss_binning_workspace_bytes_ex(lib.moss_binning_capacity(nb, ex), ex) <= nb


def test_invalid_arguments_are_rejected_without_a_gpu(lib):
    sc = _abi.MossScene()
    assert lib.moss_rasterize_forward(None, C.byref(sc), 0, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_mark_visible(None, -1, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT


def test_batch_entry_points_validate_without_a_gpu(lib):
    sc = _abi.MossScene()
    views = (_abi.MossView * 2)()
    assert lib.moss_sizeof_view() == C.sizeof(_abi.MossView)
    assert lib.moss_rasterize_forward_batch(None, C.byref(sc), 2, views, 0, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_rasterize_forward_batch(None, C.byref(sc), 2, None, 0, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    # workspace queries scale with the number of views; one view equals the single-view query
    assert lib.moss_geom_workspace_bytes_batch(20000, 1) == lib.moss_geom_workspace_bytes(20000)
    assert lib.moss_image_workspace_bytes_batch(512, 512, 1) == lib.moss_image_workspace_bytes(512, 512)
    assert lib.moss_geom_workspace_bytes_batch(20000, 23) >= 23 * 70 * 20000
    assert lib.moss_image_workspace_bytes_batch(512, 512, 23) >= 23 * 8 * 512 * 512
    assert lib.moss_ckpt_workspace_bytes_batch(100000, 512, 512, 1) == lib.moss_ckpt_workspace_bytes(100000, 512, 512)
    assert lib.moss_batch_grad_workspace_bytes(20000, 16, 1) == 0
    assert lib.moss_batch_grad_workspace_bytes(20000, 16, 23) >= 23 * 4 * 20000 * (1 + 3 + 3 + 6 + 48 + 3 + 4)


def test_scene_struct_layout_matches_header(lib):
    # ctypes mirror must have the same size as the C struct (checked through the library)
    assert lib.moss_sizeof_scene() == C.sizeof(_abi.MossScene)
    assert lib.moss_sizeof_skin() == C.sizeof(_abi.MossSkin)


def test_skin_entry_points_validate_without_a_gpu(lib):
    k = _abi.MossSkin()
    k.P, k.V = 4, 0
    assert lib.moss_skin_forward(None, C.byref(k), None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_skin_backward(None, C.byref(k), *([None] * 8)) == _abi.MOSS_ERR_INVALID_ARGUMENT
    k.P, k.V = 0, 6890       # empty input: nothing to launch, no accumulators requested
    assert lib.moss_skin_forward(None, C.byref(k), None, None, None, None, None) == _abi.MOSS_OK


def test_nearest_vertex_entry_point_validates_without_a_gpu(lib):
    assert lib.moss_nearest_vertex(None, 4, None, 0, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT    # no vertices
    assert lib.moss_nearest_vertex(None, 4, None, 10, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT   # NULL arrays
    assert lib.moss_nearest_vertex(None, 0, None, 10, None, None, None) == _abi.MOSS_OK                     # nothing to do
    from moss_b200 import skinning as K
    import torch
    with pytest.raises(RuntimeError):
        K.nearest_vertex(torch.zeros(3, 3), torch.zeros(5, 3))      # no CPU path


def test_densify_stats_entry_point_validates_without_a_gpu(lib):
    assert lib.moss_densify_stats(None, 4, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT   # no radii
    assert lib.moss_densify_stats(None, -1, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_densify_stats(None, 0, None, None, None, None, None) == _abi.MOSS_OK
    import moss_b200 as M
    with pytest.raises(RuntimeError):
        M.add_densification_stats(torch.zeros(3, dtype=torch.int32), torch.zeros(3, 3))      # no CPU path


def test_adamw_entry_point_validates_without_a_gpu(lib):
    assert lib.moss_sizeof_adam_group() == C.sizeof(_abi.MossAdamGroup)
    g = (_abi.MossAdamGroup * 2)()
    g[0].numel, g[0].lr, g[0].step = 16, 0.1, 1                 # NULL arrays
    assert lib.moss_adamw_step(None, 2, g, 0.9, 0.999, 1e-15, 0.01) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_adamw_step(None, 9, g, 0.9, 0.999, 1e-15, 0.01) == _abi.MOSS_ERR_INVALID_ARGUMENT     # > 8 groups
    g[0].numel = 0                                             # only empty groups: nothing to launch
    assert lib.moss_adamw_step(None, 2, g, 0.9, 0.999, 1e-15, 0.01) == _abi.MOSS_OK
    from moss_b200.optim import FusedAdamW
    w = torch.nn.Parameter(torch.zeros(4))
    opt = FusedAdamW([{"params": [w], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert opt.param_groups[0]["name"] == "xyz" and opt.param_groups[0]["weight_decay"] == 1e-2
    w.grad = torch.ones(4)
    with pytest.raises(RuntimeError):
        opt.step()                                             # no CPU path


def test_skinning_refuses_cpu_compute_and_frame_abi_validates():
    """Neither the per-Gaussian kernel nor the per-frame table kernel has a CPU path; argument validation of the frame
    entry points happens before any launch."""
    import ctypes as C
    from moss_b200 import skinning as K, synth
    from oracle import lbs_oracle as L
    from tests.lbs_cases import oracle_smpl
    body = synth.body()
    V = body.verts.shape[0]
    smpl = K.Smpl(body.verts, body.shapedirs, body.posedirs.t().reshape(V, 3, 207).contiguous(), body.J_regressor,
                  body.weights, body.parents)
    fr = synth.frame_params(3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        K.frame_tables(smpl, fr.params, fr.big_params)
    ft = L.frame_tables(oracle_smpl(body), fr.params, fr.big_params)
    g = synth.gaussians(64, body)
    with pytest.raises(RuntimeError, match="no CPU path"):
        K.skin_and_covariance(g.xyz, g.vert_ids, body.weights, K.FrameTables(*ft), g.scales, g.rotations)
    lib = _abi.load()
    assert lib.moss_smpl_frame_scratch_bytes(6890) == 16 + 4 * (24 * 9 + 216 * 208)
    f = _abi.MossSmplFrame()
    assert lib.moss_smpl_frame_forward(None, C.byref(f), None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_smpl_frame_backward(None, C.byref(f), None, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_smpl_joint_tables(None, 0, 10, None, None, None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT


def test_image_loss_scratch_and_tracing_switch_without_a_gpu(lib):
    """Scratch sizes of the image-loss launch (tile grid aligned to the crop: one extra tile column and row; images of up to
    296 tiles launch one block per tile and colour channel) and the NVTX switch, none of which touches the GPU."""
    tiles = lambda W, H: ((W + 31) // 32 + 1) * ((H + 31) // 32 + 1)      # noqa: E731
    assert lib.moss_image_losses_scratch_bytes(512, 512) == 16 + 16 * 3 * tiles(512, 512)       # 289 tiles: per-channel blocks
    assert lib.moss_image_losses_scratch_bytes(1920, 1080) == 16 + 16 * tiles(1920, 1080)       # 2135 tiles: one block per tile
    assert lib.moss_image_losses_scratch_bytes(0, 7) == 16
    k = _abi.MossImageLoss()
    assert lib.moss_image_losses(None, C.byref(k), None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_image_losses_static(None, C.byref(k), None, None, None) == _abi.MOSS_ERR_INVALID_ARGUMENT
    assert lib.moss_nvtx_enable(1) == 0 and lib.moss_nvtx_enable(0) == 0


def test_api_validation_and_no_cpu_fallback():
    import moss_b200 as M
    rs = M.GaussianRasterizationSettings(image_height=16, image_width=16, tanfovx=1.0, tanfovy=1.0, bg=torch.zeros(3),
                                         scale_modifier=1.0, viewmatrix=torch.eye(4), projmatrix=torch.eye(4),
                                         sh_degree=0, campos=torch.zeros(3), prefiltered=False, debug=False)
    r = M.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r(m, m, torch.zeros(4, 1), shs=None, colors_precomp=None, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    # CPU tensors: the product refuses instead of falling back to any CPU path
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))


def test_dropin_package_resolves():
    import importlib
    import sys
    p = os.path.join(ROOT, "moss_b200", "dropin")
    sys.path.insert(0, p)
    try:
        mod = importlib.import_module("diff_gaussian_rasterization")
        assert hasattr(mod, "GaussianRasterizationSettings") and hasattr(mod, "GaussianRasterizer")
        assert mod.GaussianRasterizationSettings._fields == (
            "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
            "sh_degree", "campos", "prefiltered", "debug")
    finally:
        sys.path.remove(p)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "moss_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("CPU oracle", "").replace("the oracle", "") or f == "synth.py", f


def test_compiled_autograd_binding_builds_and_loads(lib):
    """moss_b200/_C.so (csrc/torch_binding.cpp): builds against this interpreter's torch, links the C-ABI library and
    refuses CPU tensors (no CPU path)."""
    build.build_binding()
    from moss_b200 import rasterizer as R
    import moss_b200 as M
    b = R._load_binding()
    assert b is not None and b.abi_version() == lib.moss_abi_version()
    rs = M.GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3, torch.zeros(3), False, False)
    with pytest.raises(RuntimeError, match="CUDA tensors"):
        M.GaussianRasterizer(rs)(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), shs=torch.zeros(4, 16, 3),
                                 scales=torch.zeros(4, 3), rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="excatly one of either SHs"):
        M.GaussianRasterizer(rs)(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), scales=torch.zeros(4, 3), rotations=torch.zeros(4, 4))
