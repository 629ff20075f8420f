"""GPU parity of the forward path (K1-K6) through the C ABI against the CPU oracle.

Bars (BASELINE.json north_star): pre-sort pairs, sorted keys, per-tile lists, tile ranges and lastContribIdx
bit-exact; image max-abs <= 1e-4 (the pinned arithmetic actually makes it bit-exact, asserted as such).
"""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu

@pytest.fixture(params=["rounds", "warps"], autouse=True)
def forward_walk(request, gs):
    """every test of this file runs with both forward blend kernels: the CTA-cooperative rounds (the default of a render-only
    frame) and the warp-autonomous walk (gs_tune "fwd_warp" = 2; the default of a frame that records visit words)"""
    gs.tune("fwd_warp", 2 if request.param == "warps" else 0)
    yield request.param
    gs.tune("fwd_warp", 1)


FLOAT_FIELDS = ["screenPos", "conic", "depth", "opacity", "color", "radius", "viewPos_xy", "cov2D"]
UINT_FIELDS = ["tileMinX", "tileMinY", "tileMaxX", "tileMaxY"]


def run_cuda(gs, ctx, g, u, W, H):
    G = gs.Gaussians(ctx, len(g))
    G.upload(g)
    R = gs.Rasterizer(ctx, max(len(g), 64))
    img = gs.Image(ctx, W, H)
    R.forward(G, u, img)
    ctx.sync()
    return G, R, img


def check_forward(gs, ctx, orc, g, u, W, H):
    ref = orc.render(g, u)
    G, R, img = run_cuda(gs, ctx, g, u, W, H)
    try:
        proj = R.get(gs.RAST_PROJECTED)
        for f in UINT_FIELDS:
            assert np.array_equal(proj[f], ref["proj"][f]), f
        for f in FLOAT_FIELDS:
            a, b = proj[f], ref["proj"][f]
            assert np.array_equal(a, b), (f, np.abs(a - b).max())     # numeric equality (-0 == +0)
        assert R.pair_count == len(ref["keys"])
        assert np.array_equal(R.get(gs.RAST_KEYS), ref["keys"])
        assert np.array_equal(R.get(gs.RAST_VALUES), ref["vals"])
        assert np.array_equal(R.get(gs.RAST_SORTED_KEYS), ref["skeys"])
        assert np.array_equal(R.get(gs.RAST_SORTED_VALUES), ref["svals"])
        rng = R.get(gs.RAST_TILE_RANGES)
        assert np.array_equal(rng["start"], ref["ranges"]["start"])
        assert np.array_equal(rng["count"], ref["ranges"]["count"])
        assert np.array_equal(R.get(gs.RAST_LAST_IDX), ref["last"])
        out = img.download()
        assert np.abs(out - ref["image"]).max() <= 1e-4
        assert np.array_equal(out, ref["image"])                       # pinned arithmetic: bit-exact
    finally:
        img.close(); R.close(); G.close()
    return ref


@pytest.mark.parametrize("n,W,H,s0,seed", [
    (64, 64, 64, 0.05, 1),
    (1000, 64, 64, 0.03, 2),
    (1000, 100, 70, 0.03, 3),          # ragged: image not a multiple of the tile size
    (10000, 256, 256, 0.012, synth.CONFIG_SEED["C1"]),   # BASELINE config 1
    (20000, 320, 200, 0.02, 5),
])
def test_forward_parity(gs, ctx, orc, n, W, H, s0, seed):
    g = synth.make_scene(n, seed, s0)
    u = synth.make_camera(1, 8, W, H, n)
    ref = check_forward(gs, ctx, orc, g, u, W, H)
    assert len(ref["keys"]) > 0


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("n,W,H,s0,seed", [
    (3000, 200, 136, 0.05, 31),        # several coarse bins, ragged edges, splats spanning bins
    (1500, 16, 1100, 0.03, 32),        # one tile wide, 69 tiles tall (a column of 9 coarse bins)
    (1500, 1100, 16, 0.03, 33),        # one tile tall
    (40, 16, 16, 0.2, 34),             # a single tile, a single coarse bin
])
def test_forward_parity_in_every_pair_pipeline(gs, ctx, orc, mode, n, W, H, s0, seed):
    """sort_mode 0 two-level binning (default), 1 the literal 64-bit pair sort, 2 hoisted radix sort: same oracle parity"""
    gs.tune("sort_mode", mode)
    try:
        g = synth.make_scene(n, seed, s0)
        u = synth.make_camera(1, 8, W, H, n)
        check_forward(gs, ctx, orc, g, u, W, H)
    finally:
        gs.tune("sort_mode", 0)


def test_forward_empty_and_culled(gs, ctx, orc):
    """nothing visible (all behind the camera): white image, empty ranges (SURVEY note B)"""
    n, W, H = 500, 64, 48
    g = synth.make_scene(n, 9, 0.03)
    g["position"][:, :] += 100.0
    u = synth.make_camera(0, 1, W, H, n)
    ref = check_forward(gs, ctx, orc, g, u, W, H)
    assert len(ref["keys"]) == 0
    assert np.all(ref["image"] == 1.0)


def test_forward_edge_cases(gs, ctx, orc):
    """NaN / huge positions, NaN scale, zero quaternion, extreme scales & opacities, giant splats (conic-degenerate,
    > 256 tiles), elongated 20:1 clamp — the culls of tiled_shaders.metal:120-290"""
    n, W, H = 4000, 256, 256
    g = synth.make_scene(n, 11, 0.02)
    g["position"][0, 0] = np.nan
    g["position"][1, 1] = 2e6
    g["scale"][2, 2] = np.nan
    g["rotation"][3] = 0
    g["rotation"][4] = [np.nan, 0, 0, 0]
    g["scale"][5:200] = np.log(3.0)             # giant: radius clamp 512, conic magnitude < 1e-4 at 256x256
    g["scale"][200:400, 0] = 3.0                # elongated -> 20:1 clamp
    g["scale"][200:400, 1:] = -6.0
    g["scale"][400:500] = 9.0                   # beyond the +-5 clamp
    g["opacity"][500:700] = -7.0                # sigma < 0.005 -> no pairs
    g["opacity"][700:720] = np.nan
    g["sh"][720:740, 0] = np.nan
    g["opacity"][740:900] = 12.0
    u = synth.make_camera(2, 8, W, H, n)
    check_forward(gs, ctx, orc, g, u, W, H)


def test_tfinal_matches_backward_first_pass(gs, ctx, orc):
    """the T_final the forward exports == what tiledBackward's first pass (:430-460) recomputes"""
    n, W, H = 5000, 128, 96
    g = synth.make_scene(n, 21, 0.03)
    g["opacity"] += 2.0                         # make many pixels saturate (T < 1e-4)
    u = synth.make_camera(3, 8, W, H, n)
    ref = orc.render(g, u)
    _, tf = orc.backward(g, ref["proj"], ref["svals"], ref["ranges"], u, ref["last"], ref["image"], ref["image"],
                         want_tfinal=True)
    G, R, img = run_cuda(gs, ctx, g, u, W, H)
    try:
        got = R.get(gs.RAST_TFINAL)
        m = ref["last"] != 0xFFFFFFFF
        assert m.sum() > 0 and (ref["image"][..., 0][m] < 1).any()
        assert np.array_equal(got[m], tf[m])
    finally:
        img.close(); R.close(); G.close()


def test_overlap_mode_changes_nothing(gs, ctx, orc):
    """gs_rast_set_overlap: two rasterizers alternating over independent frames, the stages before the blend on a side stream;
    every frame's image and lastContribIdx stay those of the one-stream run (and of the oracle for the first view)"""
    n, W, H = 20000, 320, 200
    g = synth.make_scene(n, 77, 0.02)
    G = gs.Gaussians(ctx, n); G.upload(g)
    Rs = [gs.Rasterizer(ctx, n), gs.Rasterizer(ctx, n)]
    imgs = [gs.Image(ctx, W, H), gs.Image(ctx, W, H)]
    cams = [synth.make_camera(v, 8, W, H, n) for v in range(8)]
    try:
        want = []
        for v in range(8):
            Rs[0].forward(G, cams[v], imgs[0])
            want.append((imgs[0].download().copy(), Rs[0].get(gs.RAST_LAST_IDX).copy()))
        ref = orc.render(g, cams[0])
        assert np.array_equal(want[0][0], ref["image"]) and np.array_equal(want[0][1], ref["last"])
        for R in Rs:
            R.set_overlap(True)
        for rep in range(3):
            for v in range(0, 8, 2):                     # two frames in flight, then look at both
                Rs[0].forward(G, cams[v], imgs[0])
                Rs[1].forward(G, cams[v + 1], imgs[1])
                for k in range(2):
                    assert np.array_equal(imgs[k].download(), want[v + k][0]), (rep, v, k)
                    assert np.array_equal(Rs[k].get(gs.RAST_LAST_IDX), want[v + k][1]), (rep, v, k)
        for R in Rs:
            R.set_overlap(False)
        Rs[1].forward(G, cams[3], imgs[1])
        assert np.array_equal(imgs[1].download(), want[3][0])
    finally:
        for o in imgs + Rs + [G]:
            o.close()
