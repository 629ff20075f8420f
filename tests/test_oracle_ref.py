"""Pins the oracle (and the host layer's file formats) to the REFERENCE'S OWN OBJECT CODE for every part of the hot path
that is not Metal: oracle/_ref/libref.so is compiled from /root/reference sources where they lie (oracle/Makefile):
  parallelRadixSort                          tiled_rasterizer.mm:27-102      <-> orc.sort_pairs
  DensityController::accumulateGradients     density_control.mm:121-185      <-> orc.accumulate_gradients
  DensityController::apply                   density_control.mm:188-501      <-> orc.densify_apply
  viewMatrixFromColmap / projectionFromColmap mtl_engine.mm:637-682          <-> orc.build_uniforms, synth.build_uniforms
  exponentialLRDecay                         mtl_engine.mm:1039-1045         <-> orc_lr_decay
  wire structs                               tiled_rasterizer.hpp, ply_loader.hpp, gradients.hpp  <-> numpy dtypes / include/gs_abi.h
  load_ply / PLYExporter::exportPLY          ply_loader.cpp, ply_exporter.hpp <-> host/ply_io.hpp
The Metal kernels themselves stay unpinned (no Metal compiler or device on Linux).  No GPU needed."""
import ctypes
import numpy as np
import pytest
from gaussiansplatting_b200 import synth
from oracle import ref as R

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref.so absent and /root/reference not present to build it")


@pytest.fixture(scope="module")
def ref():
    R.lib()
    return R


def test_wire_struct_layouts_are_the_reference_headers(ref):
    from gaussiansplatting_b200.layouts import GAUSSIAN, GRADIENTS, PROJECTED, TILERANGE, UNIFORMS
    L = ref.layouts()
    assert L["Gaussian"] == GAUSSIAN.itemsize == 112 and L["GaussianGradients"] == GRADIENTS.itemsize == 112
    assert L["ProjectedGaussian"] == PROJECTED.itemsize == 88 and L["TileRange"] == TILERANGE.itemsize == 8
    assert L["TiledUniforms"] == UNIFORMS.itemsize == 240
    for dt, pre, names in ((GAUSSIAN, "Gaussian.", ("scale", "rotation", "opacity", "sh")),
                           (GRADIENTS, "GaussianGradients.", ("rotation", "sh")),
                           (PROJECTED, "ProjectedGaussian.", ("conic", "depth", "color", "radius", "tileMinX", "viewPos_xy", "cov2D")),
                           (UNIFORMS, "TiledUniforms.", ("screenSize", "cameraPos", "numTilesX"))):
        for nme in names:
            assert dt.fields[nme][1] == L[pre + nme], (pre, nme)
    assert GRADIENTS.fields["viewspace"][1] == L["GaussianGradients.viewspace_grad_x"]


@pytest.mark.parametrize("n", [0, 1, 2, 999, 1000, 1001, 4097, 100000, 1 << 20])
def test_oracle_sort_is_the_references_cpu_radix_sort(orc, ref, n):
    """below 1000 pairs the reference uses std::sort, which is not stable (SURVEY note J): distinct keys there, heavy
    duplication (ties resolved by the stable LSD passes) above"""
    rng = np.random.default_rng(n + 7)
    if n < 1000:
        keys = rng.permutation(np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15 % (1 << 40)) + np.uint64(5))
    else:
        tiles = rng.integers(0, 8160, n, dtype=np.uint64)
        depth = rng.integers(0, 1 << 10, n, dtype=np.uint64) | np.uint64(0x80000000)     # many equal (tile, depth) keys
        keys = (tiles << np.uint64(32)) | depth
        keys[: n // 8] = rng.integers(0, 1 << 63, n // 8, dtype=np.uint64)               # and full-entropy ones
    vals = np.arange(n, dtype=np.uint32)
    rk, rv = ref.sort_pairs(keys, vals)
    for th in (1, 4):
        ok, ov = orc.sort_pairs(keys, vals, th)
        assert np.array_equal(ok, rk) and np.array_equal(ov, rv), th
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(rk, keys[order]) and np.array_equal(rv, vals[order])


def _grads(n, seed):
    from oracle.orc import GRADIENTS
    rng = np.random.default_rng(seed)
    g = np.zeros(n, GRADIENTS)
    g["viewspace"] = (rng.standard_normal((n, 2)) * 1e-3).astype(np.float32)
    g["position"] = rng.standard_normal((n, 3)).astype(np.float32)
    g["viewspace"][::11] = 0                      # not counted (gradMag > 0 fails)
    g["viewspace"][3::17, 0] = np.nan             # not counted
    g["viewspace"][5::19, 1] = np.inf             # min(inf, 1) = 1: counted with the clamp value
    g["viewspace"][7::23] = 40.0                  # clamped to 1
    return g


def test_oracle_accumulate_gradients_is_the_references(orc, ref):
    n = 5003
    acc_o, cnt_o = np.zeros(n, np.float32), np.zeros(n, np.uint32)
    acc_r, cnt_r = np.zeros(n, np.float32), np.zeros(n, np.uint32)
    for step in range(3):
        g = _grads(n, 40 + step)
        orc.accumulate_gradients(g, acc_o, cnt_o)
        ref.accumulate_gradients(g, acc_r, cnt_r)
    assert np.array_equal(cnt_o, cnt_r) and cnt_r.max() == 3 and cnt_r.min() == 0
    assert np.array_equal(acc_o.view(np.uint32), acc_r.view(np.uint32))


def _densify_inputs(n, seed, s0=0.05):
    g = synth.make_scene(n, seed, s0)
    rng = np.random.default_rng(seed)
    g["opacity"][::7] = -7.0                                  # pruned: sigma < 0.005
    g["scale"][5::11] = np.log(0.5)                           # large: split; world-size prune after iteration 3000
    g["scale"][2::13, 1] = 9.0                                # beyond the +-4 clamp
    g["rotation"][4::9] *= 1.7                                # apply() uses the RAW quaternion
    counts = rng.integers(0, 4, n).astype(np.uint32)
    accum = (rng.random(n) * 1e-3 * np.maximum(counts, 1)).astype(np.float32)
    accum[::2] *= 5.0
    return g, accum, counts


@pytest.mark.parametrize("n,iteration", [(4000, 700), (4000, 3500), (4000, 400), (4000, 14999), (4000, 15000), (1, 700), (0, 700)])
def test_oracle_densify_apply_is_the_references(orc, ref, n, iteration):
    """markers, statistics, output count, layout and every kept / cloned value bit for bit; the children of a split too
    when the oracle draws from libc rand() like the reference (test hook) -- except that the reference scales the offset by
    libm expf and the oracle by its pinned polynomial exp: a last-ulp difference of the local offset, carried through the raw-
    quaternion rotation (entries up to |q|^2), held to 16 ulp of the largest offset component"""
    g, accum, counts = _densify_inputs(n, 11)
    extent, focal, W, depth = synth.scene_extent(3.0), 460.0, 512.0, 2.0 * synth.scene_extent(3.0)
    libc = ctypes.CDLL(None)
    rout, rmark, rstats = ref.densify_apply(g, accum, counts, iteration, extent, focal, W, depth, 1234)
    orc.lib().orc_set_split_rng_libc(1)
    try:
        libc.srand(1234)
        oout, omark, ostats = orc.densify_apply(g, accum, counts, iteration, extent, focal, W, depth, 0x5EED)
    finally:
        orc.lib().orc_set_split_rng_libc(0)
    if iteration >= 15000:                                     # apply() returns before touching anything (:216-220)
        assert len(rout) == n and tuple(rstats) == (0, 0, 0) and len(oout) == n
        return
    assert np.array_equal(omark, rmark) and tuple(ostats) == tuple(rstats) and len(oout) == len(rout)
    if n >= 1000:
        assert rstats[0] > 0 and (iteration <= 500 or (rstats[1] > 0 and rstats[2] > 0))
    for f in ("scale", "rotation", "opacity", "sh"):
        assert np.array_equal(oout[f].view(np.uint32), rout[f].view(np.uint32)), f
    d = np.abs(oout["position"].astype(np.float64) - rout["position"].astype(np.float64))
    exact = d == 0
    # every non-split output is bit-equal; split children differ by at most a few ulp of the offset
    src = np.repeat(np.arange(n), np.where(rmark == 1, 0, np.where(rmark == 0, 1, 2)))
    is_child = rmark[src] == 3
    assert exact[~is_child].all()
    if is_child.any():
        off = np.abs(rout["position"][is_child] - g["position"][src[is_child]]).max(axis=1, keepdims=True)
        assert np.all(d[is_child] <= 16 * 2.0 ** -23 * np.maximum(off, np.abs(rout["position"][is_child]))), float(d.max())
        assert exact[is_child].mean() > 0.5


def test_oracle_densify_cap_unmarking_is_the_references(orc, ref):
    """MAX_GAUSSIANS = 1,500,000 (density_control.mm:27): clones are un-marked first, then splits, in index order (:358-382)"""
    n = 1_000_000
    g = synth.make_scene(n, 21, 0.012)                         # below the split threshold 0.01 * extent: clones ...
    g["scale"][::3] = np.log(0.5)                              # ... and every third one a split
    counts = np.ones(n, np.uint32)
    accum = np.full(n, 1e-3, np.float32)                       # everything above the gradient threshold: clone or split
    accum[::50] = 0
    extent = synth.scene_extent(3.0)
    rout, rmark, rstats = ref.densify_apply(g, accum, counts, 700, extent, 460.0, 512.0, 2 * extent, 1)
    oout, omark, ostats = orc.densify_apply(g, accum, counts, 700, extent, 460.0, 512.0, 2 * extent, 1)
    assert len(rout) == 1_500_000 == len(oout)
    assert np.array_equal(omark, rmark) and tuple(ostats) == tuple(rstats)
    # 1.98 M wanted: the first 480 k clones in index order lose their mark, the splits keep theirs
    clones = np.flatnonzero(rmark == 2)
    assert 0 < len(clones) < 400_000 and clones.min() > 500_000 and (rmark == 3).sum() == rstats[2] > 300_000


def test_view_projection_and_uniforms_are_the_references(orc, ref):
    for i in range(6):
        q, t = synth.orbit_pose(i, 6, 3.0 + i)
        q = (q * np.float32(1.0 + 0.1 * i)).astype(np.float32)           # un-normalised quaternions go through unchanged
        W, H = 1920 - 16 * i, 1080 + 7 * i
        fx, fy, cx, cy = 0.9 * W, 0.85 * W, W / 2.0 + i, H / 2.0 - i
        V, P = ref.view_matrix(q, t), ref.projection(fx, fy, cx, cy, W, H)
        for u in (orc.build_uniforms(q, t, fx, fy, cx, cy, W, H, 10), synth.build_uniforms(q, t, fx, fy, cx, cy, W, H, 10)):
            assert np.array_equal(u["viewMatrix"][0].view(np.uint32), V.view(np.uint32))
            assert np.array_equal(u["projectionMatrix"][0].view(np.uint32), P.view(np.uint32))


def test_lr_decay_is_the_references(orc, ref):
    f = orc.lib().orc_lr_decay
    for it, mx in ((0, 30000), (1, 30000), (12345, 30000), (29999, 30000), (30000, 30000), (40000, 30000), (7, 16)):
        got = float(f(ctypes.c_float(1.6e-4), ctypes.c_float(1.6e-6), ctypes.c_uint64(it), ctypes.c_uint64(mx)))
        assert np.float32(got) == np.float32(ref.lr_decay(1.6e-4, 1.6e-6, it, mx)), (it, mx)


def test_ply_files_cross_load_with_the_references_reader_and_writer(ref, tmp_path):
    """host/ply_io.hpp writes what the reference's load_ply reads back identically, and reads what the reference's
    PLYExporter writes; both writers produce the same bytes"""
    from gaussiansplatting_b200 import host
    g = synth.make_scene(700, 5, 0.05)
    rng = np.random.default_rng(5)
    g["sh"] = rng.standard_normal((700, 12)).astype(np.float32)
    g["position"][5, 0] = np.nan; g["position"][6, 0] = 2e6
    a, b = tmp_path / "ours.ply", tmp_path / "ref.ply"
    host.ply_export(a, g)
    ref.ply_export(b, g)
    assert open(a, "rb").read() == open(b, "rb").read()
    ours_by_ref = ref.ply_load(a)
    ref_by_ours, info = host.ply_load(b)
    ref_by_ref = ref.ply_load(b)
    assert len(ours_by_ref) == len(ref_by_ours) == len(ref_by_ref) == 698 == info["vertices"] - info["skipped"]
    for f in ("position", "scale", "rotation", "opacity", "sh"):
        assert np.array_equal(ours_by_ref[f].view(np.uint32), ref_by_ref[f].view(np.uint32)), f
        assert np.array_equal(ref_by_ours[f].view(np.uint32), ref_by_ref[f].view(np.uint32)), f
    # linear-scale files (detectLinearScales, ply_loader.cpp:17-52): scales stored as exp(log-scale)
    gl = g.copy(); gl["scale"] = np.exp(g["scale"]); gl["position"][[5, 6], 0] = 0
    ref.ply_export(b, gl)
    x, info = host.ply_load(b); y = ref.ply_load(b)
    assert info["linear_scales"]
    for f in ("position", "scale", "rotation", "opacity", "sh"):
        assert np.array_equal(x[f].view(np.uint32), y[f].view(np.uint32)), f
