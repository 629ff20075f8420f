"""The viewer draw (vertexShader / fragmentShader + blend state of MTLEngine::render, shaders.metal:160-317,
mtl_engine.mm:354-362,575-605) as CUDA compute, against the CPU restatement in oracle/viewer_oracle.cpp.

Bar: the per-instance records, the per-tile instance lists (tile ids, Gaussian ids, ranges) and the fragment counts are
bit-exact; the image is bit-exact in practice and asserted to max-abs 1e-4 (north_star's fp32 image tolerance)."""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth
from gaussiansplatting_b200.layouts import view_uniforms

pytestmark = pytest.mark.gpu


def _draw(gs, ctx, g, u, flags=0, order=None):
    n = len(g)
    W, H = int(u["screenSize"][0][0]), int(u["screenSize"][0][1])
    vu = view_uniforms(u)
    G = gs.Gaussians(ctx, max(n, 1)); G.upload(g)
    S = gs.Sort32(ctx, max(n, 1))
    V = gs.Viewer(ctx, max(n, 1))
    img = gs.Image(ctx, W, H)
    pos = ctx.alloc(max(n, 1) * 16)
    try:
        if order is None:
            G.positions(pos)
            S.sort_depth(pos, u["cameraPos"][0][:3], n)
            d_order = S.sorted_indices
            host_order = ctx.download(d_order, (n,), np.uint32)
        else:
            host_order = np.ascontiguousarray(order, np.uint32)
            ctx.upload(pos, host_order)
            d_order = pos
        V.render(G, d_order, vu, img, flags)                  # the product path: patches a splat cannot reach are skipped
        fast = img.download()
        V.render(G, d_order, vu, img, flags | gs.GS_VIEW_COUNT_FRAGMENTS)   # counting path: every quad fragment is shaded
        assert np.array_equal(fast.view(np.uint32), img.download().view(np.uint32)), "footprint skipping changed the image"
        ntx, nty = (W + 15) // 16, (H + 15) // 16
        return dict(image=fast, order=host_order, splats=V.get(gs.VIEW_SPLATS, n=n), tiles=V.get(gs.VIEW_SORTED_TILES),
                    vals=V.get(gs.VIEW_SORTED_VALUES), ranges=V.get(gs.VIEW_TILE_RANGES, tiles=ntx * nty), frags=V.fragments(),
                    pairs=V.pair_count)
    finally:
        ctx.free(pos)
        for o in (img, V, S, G):
            o.close()


def _expected_lists(sp, order, ntx, nty):
    """per-tile instance lists in draw order from the oracle's per-instance records"""
    lists = [[] for _ in range(ntx * nty)]
    for i in order:
        s = sp[i]
        if not s["valid"]:
            continue
        for ty in range(s["tileMinY"], s["tileMaxY"] + 1):
            for tx in range(s["tileMinX"], s["tileMaxX"] + 1):
                lists[ty * ntx + tx].append(i)
    return lists


def _check(gs, ctx, orc, g, u, quantize=False):
    vu = orc.view_uniforms(u)
    W, H = int(u["screenSize"][0][0]), int(u["screenSize"][0][1])
    got = _draw(gs, ctx, g, u, gs.GS_VIEW_QUANTIZE_RGBA8 if quantize else 0)
    want_order = orc.viewer_order(g, u["cameraPos"][0][:3])
    assert np.array_equal(got["order"], want_order), "draw order (GPURadixSort32) differs from the oracle"
    sp = orc.viewer_splats(g, vu)
    for f in sp.dtype.names:
        if f != "_pad":
            assert np.array_equal(got["splats"][f].view(np.uint32), sp[f].view(np.uint32)), f
    ntx, nty = (W + 15) // 16, (H + 15) // 16
    lists = _expected_lists(sp, want_order, ntx, nty)
    assert got["pairs"] == sum(len(l) for l in lists)
    start = 0
    for t, l in enumerate(lists):
        assert got["ranges"]["count"][t] == len(l), t
        if l:
            assert got["ranges"]["start"][t] == start
            assert np.array_equal(got["vals"][start:start + len(l)], np.asarray(l, np.uint32)), t
            assert np.all(got["tiles"][start:start + len(l)] == t)
        start += len(l)
    img, shaded, blended = orc.viewer_render(g, want_order, vu, quantize=quantize, threads=0)
    assert got["frags"] == (shaded, blended), "fragment counts differ from the oracle"
    assert np.abs(got["image"] - img).max() <= 1e-4
    return got, img


@pytest.mark.parametrize("n,W,H,scale,seed", [(2000, 128, 96, 0.03, 123), (10000, 256, 256, 0.012, 7), (500, 100, 70, 0.2, 9),
                                              (300, 16, 16, 0.05, 3), (4000, 333, 17, 0.02, 11)])
def test_viewer_parity(gs, ctx, orc, n, W, H, scale, seed):
    g = synth.make_scene(n, seed, scale)
    u = synth.make_camera(seed % 5, 8, W, H, n)
    got, img = _check(gs, ctx, orc, g, u)
    assert img.min() < 0.99, "the scene must actually draw something"


def test_viewer_rgba8_target(gs, ctx, orc):
    g = synth.make_scene(3000, 21, 0.03)
    u = synth.make_camera(2, 8, 160, 120, 3000)
    got, img = _check(gs, ctx, orc, g, u, quantize=True)
    assert np.abs(np.rint(img * 255.0) - img * 255.0).max() < 1e-3, "an RGBA8Unorm target holds multiples of 1/255"


def test_viewer_edge_cases(gs, ctx, orc):
    n, W, H = 1500, 96, 80
    g = synth.make_scene(n, 5, 0.04)
    rng = np.random.default_rng(0)
    g["position"][:50] = np.nan                               # vertexShader's invalid-position path
    g["position"][50:100] *= 1e7                              # |position| > 1e6
    g["position"][100:150, 2] += 50.0                         # far behind / beyond the camera
    g["scale"][150:200] = 9.0                                 # beyond the +-8 clamp: radius clamp at 512, conic rescale
    g["scale"][200:250] = -12.0                               # tiny: the 0.3 floor
    g["rotation"][250:300] *= 3.0                             # raw (un-normalised) quaternions scale the covariance
    g["rotation"][300:320] = 0.0
    g["opacity"][320:360] = -9.0                              # alpha below 1/255 everywhere
    g["position"][360:400] = rng.uniform(-0.05, 0.05, (40, 3)).astype(np.float32) + np.array([0, 0, 0], np.float32)
    u = synth.make_camera(1, 8, W, H, n)
    _check(gs, ctx, orc, g, u)


def test_viewer_empty_scene_is_the_clear_colour(gs, ctx, orc):
    g = synth.make_scene(64, 1, 0.01)
    g["position"][:] = np.nan
    u = synth.make_camera(0, 8, 48, 32, 64)
    got = _draw(gs, ctx, g, u)
    assert got["pairs"] == 0 and np.all(got["image"] == 1.0)


def test_viewer_order_is_the_callers(gs, ctx, orc):
    """the draw follows the index buffer it is given (instance j = sortedIndices[j], mtl_engine.mm:597): reversed order
    gives the oracle's reversed-order image, which differs from the back-to-front one"""
    n, W, H = 1500, 96, 64
    g = synth.make_scene(n, 77, 0.05)
    u = synth.make_camera(3, 8, W, H, n)
    vu = orc.view_uniforms(u)
    order = orc.viewer_order(g, u["cameraPos"][0][:3])
    rev = order[::-1].copy()
    got = _draw(gs, ctx, g, u, order=rev)
    want, _, _ = orc.viewer_render(g, rev, vu)
    base, _, _ = orc.viewer_render(g, order, vu)
    assert np.abs(got["image"] - want).max() <= 1e-4
    assert np.abs(want - base).max() > 1e-2


def test_viewer_rejects_a_mismatched_target(gs, ctx):
    g = synth.make_scene(64, 1, 0.01)
    u = synth.make_camera(0, 8, 48, 32, 64)
    vu = view_uniforms(u)
    G = gs.Gaussians(ctx, 64); G.upload(g)
    V = gs.Viewer(ctx, 64)
    img = gs.Image(ctx, 64, 32)
    pos = ctx.alloc(64 * 4)
    try:
        with pytest.raises(gs.GsError):
            V.render(G, pos, vu, img)
    finally:
        ctx.free(pos)
        for o in (img, V, G):
            o.close()
