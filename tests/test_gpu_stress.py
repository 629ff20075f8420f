"""Repeatability stress of the kernels built on spin-waits and tickets (decoupled look-back in the scans and the onesweep
passes, the self-resetting tickets of bin_scan / the tile-order CTA, the side-stream fork / join of the bin map): hundreds of
frames and sorts whose every integer output must be bit-identical to the first run / to numpy.  compute-sanitizer is closed on
this GPU pool (profiles/r2_sanitizer_unavailable.txt); a protocol race shows up here as a mismatch or a hang (pytest-timeout)."""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.timeout(300)
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_many_frames_give_identical_lists_and_images(gs, ctx, mode):
    n, W, H = 30000, 400, 240
    G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, 211, 0.03))
    R = gs.Rasterizer(ctx, n); img = gs.Image(ctx, W, H)
    cams = [synth.make_camera(v, 5, W, H, n) for v in range(5)]
    gs.tune("sort_mode", mode)
    try:
        ref = []
        for u in cams:
            R.forward(G, u, img)
            ref.append((R.get(gs.RAST_SORTED_VALUES), R.get(gs.RAST_TILE_RANGES), R.get(gs.RAST_LAST_IDX), img.download()))
        assert min(len(r[0]) for r in ref) > 50000
        for it in range(120):
            v = (it * 3) % 5
            R.forward(G, cams[v], img)
            if it % 4 == 0:                        # every frame is rendered, every fourth is read back and compared
                sv, rg, last, im = ref[v]
                assert np.array_equal(R.get(gs.RAST_SORTED_VALUES), sv), (mode, it)
                got = R.get(gs.RAST_TILE_RANGES)
                assert np.array_equal(got["start"], rg["start"]) and np.array_equal(got["count"], rg["count"]), (mode, it)
                assert np.array_equal(R.get(gs.RAST_LAST_IDX), last) and np.array_equal(img.download(), im), (mode, it)
    finally:
        gs.tune("sort_mode", 0)
        img.close(); R.close(); G.close()


@pytest.mark.timeout(300)
def test_many_sorts_of_many_sizes_match_numpy(gs, ctx):
    rng = np.random.default_rng(5)
    S = gs.Sort64(ctx, 1 << 21)
    try:
        for it in range(60):
            n = int(rng.integers(1, 1 << 21)) if it % 3 else int(rng.choice([1, 255, 256, 257, 4607, 4608, 4609, 1 << 20, (1 << 20) + 1]))
            kind = it % 4
            if kind == 0:
                keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
            elif kind == 1:
                keys = rng.integers(0, 1 << 12, n, dtype=np.uint64) << np.uint64(rng.integers(0, 50))     # few distinct values
            elif kind == 2:
                keys = np.sort(rng.integers(0, 1 << 40, n, dtype=np.uint64))[::-1].copy()                # reverse sorted
            else:
                keys = (rng.integers(0, 8160, n, dtype=np.uint64) << np.uint64(32)) | rng.integers(1 << 31, 1 << 32, n, dtype=np.uint64)
            vals = np.arange(n, dtype=np.uint32)
            k, v = S.sort_host(keys, vals)
            o = np.argsort(keys, kind="stable")
            assert np.array_equal(k, keys[o]) and np.array_equal(v, vals[o]), (it, n, kind)
    finally:
        S.close()


@pytest.mark.timeout(300)
def test_backward_moment_buffer_stays_clean_over_many_frames(gs, ctx):
    """K8 clears the moment rows K7 wrote (no per-frame memset): after many fwd+bwd frames over changing views the gradients of a
    view must still equal its first-run gradients to the atomic-order tolerance, and untouched Gaussians stay exactly zero"""
    n, W, H = 20000, 320, 200
    G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, 311, 0.03))
    R = gs.Rasterizer(ctx, n); D = gs.Grads(ctx, n)
    img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    gti.upload(np.full((H, W, 4), 0.4, np.float32))
    cams = [synth.make_camera(v, 4, W, H, n) for v in range(4)]
    try:
        first = []
        for u in cams:
            R.forward(G, u, img); R.backward(G, D, u, img, gti)
            first.append(D.download(n))
        for it in range(60):
            v = (it * 3) % 4
            R.forward(G, cams[v], img); R.backward(G, D, cams[v], img, gti)
            if it % 6 == 0:
                got = D.download(n)
                for f in ("position", "opacity", "scale", "rotation", "sh"):
                    a, b = got[f].astype(np.float64), first[v][f].astype(np.float64)
                    assert np.array_equal(a == 0, b == 0), (f, it)                   # the same Gaussians are touched
                    scale = np.abs(b).max()
                    assert np.all(np.abs(a - b) <= 1e-3 * np.maximum(np.abs(a), np.abs(b)) + 1e-6 * scale), (f, it)
    finally:
        for o in (gti, img, D, R, G):
            o.close()
