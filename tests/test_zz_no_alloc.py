"""gs_loss_compute_noalloc (include/gs_abi.h): the logging loss with its scratch in a static device array and no allocation,
for callers that cannot call cudaMalloc.  No counterpart in the reference.
Checked through the C ABI against the existing entry point gs_loss_compute on the same images, at the tolerance
tests/test_gpu_loss.py uses for this scalar (2e-6 absolute), on calls in a row that share the static scratch: a large image
first, then smaller ones, so stale partial sums of an earlier call would show.  Collected last on purpose (pytest -x)."""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu
TOL = 2e-6


def _images(gs, ctx, W, H, seed):
    rng = np.random.default_rng(seed)
    a = rng.random((H, W, 4), dtype=np.float32); b = np.clip(a + 0.2 * rng.standard_normal((H, W, 4)).astype(np.float32), 0, 1)
    a[..., 3] = 1; b[..., 3] = 1
    ia, ib = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    ia.upload(a); ib.upload(b)
    return ia, ib


def test_noalloc_loss_matches_the_allocating_entry_point_on_calls_in_a_row(gs, ctx):
    d_out = ctx.alloc(4)                                  # the caller's output: one float of device memory
    try:
        seen = []
        for k, (W, H, quant) in enumerate([(640, 360, False), (100, 70, False), (256, 256, True), (100, 70, False), (16, 16, False)]):
            ia, ib = _images(gs, ctx, W, H, 100 + k)
            try:
                want = gs.loss(ctx, ia, ib, 0.2, quant)                       # existing entry point (allocates its accumulator)
                ctx.upload(d_out, np.array([np.nan], np.float32))
                gs.loss_noalloc(ctx, ia, ib, d_out, 0.2, quant)               # enqueues only
                ctx.sync()
                got = float(ctx.download(d_out, 1, np.float32)[0])
                gs.loss_noalloc(ctx, ia, ib, d_out, 0.2, quant)               # again: fixed-order sum, so bit-identical
                ctx.sync()
                again = float(ctx.download(d_out, 1, np.float32)[0])
            finally:
                ia.close(); ib.close()
            assert np.isfinite(got) and 0.01 < want < 1.0
            assert abs(got - want) <= TOL, (W, H, got, want)
            assert again == got
            seen.append(got)
        assert abs(seen[1] - seen[3]) > 1e-5              # different images (seeds), so equal answers would mean stale scratch
    finally:
        ctx.free(d_out)


def test_noalloc_loss_on_a_rendered_frame(gs, ctx):
    n, W, H = 3000, 200, 120
    G = gs.Gaussians(ctx, n); R = gs.Rasterizer(ctx, n)
    img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    d_out = ctx.alloc(4)
    try:
        u = synth.make_camera(1, 8, W, H, n)
        G.upload(synth.make_scene(n, 9 ^ 0xA5A5, 0.04)); R.forward(G, u, gti)
        G.upload(synth.make_scene(n, 9, 0.04)); R.forward(G, u, img)
        want = gs.loss(ctx, img, gti)
        gs.loss_noalloc(ctx, img, gti, d_out); ctx.sync()
        assert abs(float(ctx.download(d_out, 1, np.float32)[0]) - want) <= TOL
    finally:
        ctx.free(d_out)
        for o in (gti, img, R, G):
            o.close()


def test_noalloc_loss_rejects_what_the_static_scratch_cannot_hold(gs, ctx):
    W, H = 4096, 2064                                     # 256 x 129 = 33,024 tiles > GS_LOSS_NOALLOC_MAX_TILES = 32,768
    ia, ib = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    small, other = gs.Image(ctx, 32, 32), gs.Image(ctx, 48, 32)
    d_out = ctx.alloc(4)
    try:
        with pytest.raises(gs.GsError, match="exceeds the static scratch"):
            gs.loss_noalloc(ctx, ia, ib, d_out)
        with pytest.raises(gs.GsError, match="sizes differ"):
            gs.loss_noalloc(ctx, small, other, d_out)
        with pytest.raises(gs.GsError, match="null"):
            gs.loss_noalloc(ctx, small, small, None)
    finally:
        ctx.free(d_out)
        for o in (other, small, ib, ia):
            o.close()
