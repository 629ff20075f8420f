"""GPU parity of the backward path (K7 + K8) through the C ABI against the CPU oracle.

Bar (north_star): gradients within 1e-3 relative, "since atomic ordering differs".  Per component, against the oracle's
double-accumulated value b:   |a-b| <= 1e-3 * max(|a|,|b|) + k * 2^-23 * sum|term|   where k and sum|term| are the
component's OWN number of accumulated (pixel, Gaussian) terms and the sum of their magnitudes, both produced by the
oracle (orc.backward(want_abs=True)): the band a float-atomic accumulation in arbitrary order (the reference,
tiled_shaders.metal:699-736) can move in.  There is no group-wide floor.  oracle/parity.py holds the rule and reports, per
parameter group, the worst PURE relative error and the fraction of components that need the band.
"""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth
from oracle import parity

pytestmark = pytest.mark.gpu

GROUPS = list(parity.GROUPS)


def compare_grads(got, ref, absref, count):
    """components the oracle itself overflows on (inf/NaN: giant splats, denom^2 > FLT_MAX) must be non-finite on
    the GPU too; every finite component is held to the tolerance"""
    rep = parity.assert_grads(got, ref, absref, count)
    print({f: (r["worst_pure_rel"], r["frac_over_rtol_pure"], r["frac_need_band"]) for f, r in rep.items()})
    return rep


@pytest.fixture(params=["pinned", "visits", "visits-rounds"], autouse=True)
def walk(request, gs):
    """every test of this file runs twice: with the backward walk that re-derives its decisions (gs_tune "bwd_masks" = 0; also the
    only one a rasterizer's FIRST backward can take), with the walk on the visit words the forward recorded (every later frame
    of a training caller; warp-autonomous kernels) and with the CTA-cooperative kernels on such words"""
    gs.tune("bwd_masks", 0 if request.param == "pinned" else 1)
    rounds = request.param == "visits-rounds"       # the CTA-cooperative pair of kernels instead of the warp-autonomous one
    gs.tune("fwd_warp", 0 if rounds else 1); gs.tune("bwd_warp", 0 if rounds else 1)
    yield request.param
    gs.tune("bwd_masks", 1); gs.tune("fwd_warp", 1); gs.tune("bwd_warp", 1)


def run_backward(gs, R, G, D, u, img, gti, flags, walk):
    """forward + backward in the requested walk; the visits walk needs one earlier backward on the rasterizer (it allocates the
    words) and is asserted to have been taken"""
    if walk != "pinned":
        R.forward(G, u, img)
        R.backward(G, D, u, img, gti, flags)
    R.forward(G, u, img)
    R.backward(G, D, u, img, gti, flags)
    assert R.backward_used_visits == (walk != "pinned")


def fwd_bwd(gs, ctx, orc, g, gt_g, u, W, H, quantize=False, walk="pinned"):
    ref = orc.render(g, u)
    gt = orc.render(gt_g, u)["image"]
    ref_gr = orc.backward(g, ref["proj"], ref["svals"], ref["ranges"], u, ref["last"], ref["image"], gt, quantize=quantize, want_abs=True)
    G = gs.Gaussians(ctx, len(g)); G.upload(g)
    R = gs.Rasterizer(ctx, len(g))
    img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    gti.upload(gt)
    D = gs.Grads(ctx, len(g))
    try:
        run_backward(gs, R, G, D, u, img, gti, gs.GS_BWD_QUANTIZE_RGBA8 if quantize else 0, walk)
        got = D.download(len(g))
    finally:
        D.close(); gti.close(); img.close(); R.close(); G.close()
    return got, ref_gr


@pytest.mark.parametrize("n,W,H,s0,seed", [
    (300, 64, 64, 0.05, 7),
    (2000, 100, 70, 0.03, 8),
    (10000, 256, 256, 0.012, synth.CONFIG_SEED["C1"]),
])
def test_backward_parity(gs, ctx, orc, n, W, H, s0, seed, walk):
    g = synth.make_scene(n, seed, s0)
    gt_g = synth.make_scene(n, seed ^ 0xA5A5, s0)
    u = synth.make_camera(1, 8, W, H, n)
    got, (ref, ab, cnt) = fwd_bwd(gs, ctx, orc, g, gt_g, u, W, H, walk=walk)
    assert np.abs(ref["opacity"]).max() > 0
    rep = compare_grads(got, ref, ab, cnt)
    # almost every component is inside the PURE 1e-3 relative bar; the band is for the cancelling remainder
    assert all(r["frac_need_band"] < 0.02 for r in rep.values()), rep
    # SH-rest gradients are never written (tiled_shaders.metal:699-704)
    for k in (1, 2, 3, 5, 6, 7, 9, 10, 11):
        assert np.all(got["sh"][:, k] == 0)


def test_backward_quantized_and_saturated(gs, ctx, orc, walk):
    """RGBA8 render-target mode (note G) on a dense scene where many pixels hit the T < 1e-4 cut (note I)"""
    n, W, H = 6000, 128, 96
    g = synth.make_scene(n, 31, 0.03); g["opacity"] += 2.0
    gt_g = synth.make_scene(n, 32, 0.03)
    u = synth.make_camera(5, 8, W, H, n)
    got, (ref, ab, cnt) = fwd_bwd(gs, ctx, orc, g, gt_g, u, W, H, quantize=True, walk=walk)
    compare_grads(got, ref, ab, cnt)


def test_backward_degenerate_conic_tiles(gs, ctx, orc, walk):
    """giant splats whose conic magnitude is < 1e-4 are skipped by the forward but not by the backward's first
    pass (note I): exercises the per-tile recompute path"""
    n, W, H = 3000, 256, 256
    g = synth.make_scene(n, 41, 0.02)
    g["scale"][:6] = np.log(3.0)            # six faint giants: enough to flag every tile, not enough to blow T up
    g["opacity"][:6] = -2.5
    gt_g = synth.make_scene(n, 42, 0.02)
    u = synth.make_camera(2, 8, W, H, n)
    ref = orc.render(g, u)
    p = ref["proj"]
    mag = np.abs(p["conic"]).sum(1)
    assert ((mag < 1e-4) & (p["radius"] > 0)).sum() > 0, "fixture must contain conic-degenerate visible splats"
    got, (refg, ab, cnt) = fwd_bwd(gs, ctx, orc, g, gt_g, u, W, H, walk=walk)
    assert np.isfinite(refg["scale"]).all()
    compare_grads(got, refg, ab, cnt)


def test_backward_nonfinite_pathology(gs, ctx, orc, walk):
    """150 opaque giants: the reference's second pass divides T by (1-alpha) for entries its first pass never
    multiplied in (note I), T overflows and the gradients of ~1/3 of the scene become inf/NaN.  The CUDA path must
    be non-finite for (at least) the same Gaussians so that Adam's NaN guard (shaders.metal:567-570) skips them."""
    n, W, H = 3000, 256, 256
    g = synth.make_scene(n, 41, 0.02)
    g["scale"][:150] = np.log(3.0)
    gt_g = synth.make_scene(n, 42, 0.02)
    u = synth.make_camera(2, 8, W, H, n)
    got, (refg, _, _) = fwd_bwd(gs, ctx, orc, g, gt_g, u, W, H, walk=walk)
    bad_ref = ~(np.isfinite(refg["position"][:, 0]) & np.isfinite(refg["opacity"]) & np.isfinite(refg["sh"][:, 0]))
    bad_got = ~(np.isfinite(got["position"][:, 0]) & np.isfinite(got["opacity"]) & np.isfinite(got["sh"][:, 0]))
    assert bad_ref.sum() > 100
    assert np.all(bad_got[bad_ref])


def test_backward_accumulate_two_views(gs, ctx, orc):
    """GS_BWD_ACCUMULATE sums views (the data-parallel step) and per-view stat mode feeds density control"""
    n, W, H = 2000, 96, 64
    g = synth.make_scene(n, 51, 0.04)
    gt_g = synth.make_scene(n, 52, 0.04)
    G = gs.Gaussians(ctx, n); G.upload(g)
    R = gs.Rasterizer(ctx, n)
    img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    D = gs.Grads(ctx, n); D.set_stat_mode(True)
    total = tabs = tcnt = None
    accum, count = np.zeros(n, np.float32), np.zeros(n, np.uint32)
    try:
        for v in range(2):
            u = synth.make_camera(v, 2, W, H, n)
            ref = orc.render(g, u)
            gt = orc.render(gt_g, u)["image"]
            rg, ab, cn = orc.backward(g, ref["proj"], ref["svals"], ref["ranges"], u, ref["last"], ref["image"], gt, want_abs=True)
            orc.accumulate_gradients(rg, accum, count)
            if v == 0:
                total, tabs, tcnt = rg.copy(), ab.copy(), cn.copy()
            else:
                for f in GROUPS:
                    total[f] = total[f] + rg[f]
                    tabs[f] = tabs[f] + ab[f]
                tcnt = tcnt + cn
            gti.upload(gt)
            R.forward(G, u, img)
            R.backward(G, D, u, img, gti, gs.GS_BWD_ACCUMULATE if v else 0)
        got = D.download(n)
        parity.assert_grads(got, total, tabs, tcnt, groups=GROUPS[:-1])
        # columns 14/15 = sum of per-view clamped norms, number of views counted
        assert np.array_equal(got["viewspace"][:, 1].astype(np.uint32), count)
        assert np.allclose(got["viewspace"][:, 0], accum, rtol=1e-3, atol=1e-6)
    finally:
        D.close(); gti.close(); img.close(); R.close(); G.close()
