"""Asynchronous RGBA8 image transfers (gs_image_upload_rgba8_async / download_rgba8_async / wait): a pipelined loop that
prefetches the next ground truth and reads the last render back while the blend kernels run must produce exactly what the
synchronous calls produce."""
import numpy as np
import pytest
import torch
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu


def test_pipelined_transfers_equal_synchronous_ones(gs, ctx):
    n, W, H = 20000, 320, 200
    G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, 11, 0.03))
    R = gs.Rasterizer(ctx, n)
    D = gs.Grads(ctx, n)
    cams = [synth.make_camera(v, 8, W, H, n, 3.0) for v in range(6)]
    rng = np.random.default_rng(5)
    gts = [rng.integers(0, 256, (H, W, 4), dtype=np.uint8) for _ in cams]
    img, gt = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    # synchronous reference
    want_img, want_grad = [], []
    for u, t in zip(cams, gts):
        gt.upload_rgba8(t)
        R.forward(G, u, img); R.backward(G, D, u, img, gt)
        want_img.append(img.download_rgba8()); want_grad.append(D.download(n))
    # pipelined: two images each way, pinned buffers
    imgs, gtis = [gs.Image(ctx, W, H) for _ in range(2)], [gs.Image(ctx, W, H) for _ in range(2)]
    pin_in = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in cams]
    pin_out = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in cams]
    for p, t in zip(pin_in, gts):
        p.numpy()[...] = t
    gtis[0].upload_rgba8_async(pin_in[0].data_ptr())
    for i, u in enumerate(cams):
        if i + 1 < len(cams):
            gtis[(i + 1) % 2].upload_rgba8_async(pin_in[i + 1].data_ptr())
        R.forward(G, u, imgs[i % 2])
        imgs[i % 2].download_rgba8_async(pin_out[i].data_ptr())
        R.backward(G, D, u, imgs[i % 2], gtis[i % 2])
        got = D.download(n)
        for f in ("position", "opacity", "scale", "rotation", "sh"):
            a, b = got[f], want_grad[i][f]
            assert np.allclose(a, b, rtol=1e-3, atol=1e-6 * max(1e-30, float(np.abs(b).max())), equal_nan=True), (i, f)
        if i >= 1:
            imgs[(i - 1) % 2].wait()
            assert np.array_equal(pin_out[i - 1].numpy(), want_img[i - 1]), i - 1
    imgs[(len(cams) - 1) % 2].wait()
    assert np.array_equal(pin_out[-1].numpy(), want_img[-1])
    for o in imgs + gtis + [img, gt, D, R, G]:
        o.close()
