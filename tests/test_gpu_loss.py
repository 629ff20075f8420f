"""gs_loss_compute (MTLEngine::computeLoss: L1 + D-SSIM logging loss, mtl_engine.mm:769-853) against the oracle.
Tolerance 2e-6 absolute on the scalar: the kernel keeps the reference's two-pass window arithmetic in fp32 but sums the
pixels in double, and CUDA's expf differs from libm's in the last ulp."""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("W,H", [(64, 48), (100, 70), (7, 5), (16, 16), (333, 129)])
def test_loss_parity(gs, ctx, orc, W, H):
    rng = np.random.default_rng(W * 1000 + H)
    a = rng.random((H, W, 4)).astype(np.float32)
    b = np.clip(a + 0.3 * rng.standard_normal((H, W, 4)), 0, 1).astype(np.float32)
    A, B = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    try:
        A.upload(a); B.upload(b)
        for lam in (0.2, 0.0, 1.0):
            assert abs(gs.loss(ctx, A, B, lam) - orc.loss(a, b, lam)) <= 2e-6, lam
        assert abs(gs.loss(ctx, A, B, 0.2, quantize=True) - orc.loss(a, b, 0.2, quantize=True)) <= 2e-6
        assert gs.loss(ctx, A, A) <= 1e-7
    finally:
        A.close(); B.close()


def test_loss_of_a_rendered_frame(gs, ctx, orc):
    n, W, H = 5000, 160, 128
    G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, 5, 0.03))
    Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, 6, 0.03))
    R = gs.Rasterizer(ctx, n)
    img, gt = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    try:
        u = synth.make_camera(1, 8, W, H, n)
        R.forward(G, u, img); R.forward(Gt, u, gt)
        got = gs.loss(ctx, img, gt, 0.2, quantize=True)
        want = orc.loss(img.download(), gt.download(), 0.2, quantize=True)
        assert 0.0 < want < 1.0 and abs(got - want) <= 2e-6
    finally:
        for o in (img, gt, R, G, Gt):
            o.close()
