"""GPU parity against the committed fixtures in tests/golden/ (dumped by the single-thread oracle, see
tests/golden/make_golden.py): the CUDA path through the C ABI must reproduce every intermediate without the oracle
being present.  Integers bit-exact; image <= 1e-4; gradients <= 1e-3 relative; Adam / densify layout exact."""
import glob
import os
import numpy as np
import pytest
from gaussiansplatting_b200 import synth
from test_gpu_backward import compare_grads

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
LRS = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_reproduces_golden(gs, ctx, path):
    d = np.load(path)
    g, u = d["gaussians"], d["uniforms"]
    n = len(g)
    W, H = int(u["screenSize"][0][0]), int(u["screenSize"][0][1])
    G = gs.Gaussians(ctx, n); G.upload(g)
    R = gs.Rasterizer(ctx, n)
    img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    gti.upload(d["gt"])
    D = gs.Grads(ctx, n); A = gs.Adam(ctx, n); Dn = gs.Density(ctx)
    try:
        R.forward(G, u, img)
        proj = R.get(gs.RAST_PROJECTED)
        for f in proj.dtype.names:
            if not f.startswith("_pad"):
                assert np.array_equal(proj[f], d["proj"][f]), f
        assert np.array_equal(R.get(gs.RAST_KEYS), d["keys"]) and np.array_equal(R.get(gs.RAST_VALUES), d["vals"])
        assert np.array_equal(R.get(gs.RAST_SORTED_KEYS), d["skeys"])
        assert np.array_equal(R.get(gs.RAST_SORTED_VALUES), d["svals"])
        rng = R.get(gs.RAST_TILE_RANGES)
        assert np.array_equal(rng["start"], d["ranges"]["start"]) and np.array_equal(rng["count"], d["ranges"]["count"])
        assert np.array_equal(R.get(gs.RAST_LAST_IDX), d["last"])
        out = img.download()
        assert np.abs(out - d["image"]).max() <= 1e-4
        assert R.count_evals() == tuple(int(x) for x in d["evals"])
        R.forward(G, u, img)
        R.backward(G, D, u, img, gti)
        got = D.download(n)
        compare_grads(got, d["grads"], d["grads_abs"], d["grads_terms"])
        assert not R.backward_used_visits            # a rasterizer's first backward re-derives the walk's decisions ...
        R.forward(G, u, img)
        R.backward(G, D, u, img, gti)
        assert R.backward_used_visits                # ... every later one takes them from the forward's visit words
        assert np.array_equal(R.get(gs.RAST_LAST_IDX), d["last"]) and np.abs(img.download() - d["image"]).max() <= 1e-4
        compare_grads(D.download(n), d["grads"], d["grads_abs"], d["grads_terms"])
        # Adam on the ORACLE's gradients (so the comparison is exact, not 1e-3)
        D.upload(d["grads"])
        A.step(G, D, LRS)
        after = G.download()
        for f in ("position", "scale", "rotation", "opacity", "sh"):
            assert np.array_equal(after[f], d["adam_gaussians"][f]), f
        Dn.accumulate(D, n)
        acc, cnt = Dn.download(n)
        assert np.array_equal(acc, d["dens_accum"]) and np.array_equal(cnt, d["dens_counts"])
        # densify on the fixture's inputs
        G.upload(d["dens_in"])
        Dn.set_scene_extent(synth.scene_extent(3.0)); Dn.set_seed(0x5EED)
        Dn.upload(d["dens_accum_in"], d["dens_counts"])
        stats = Dn.apply(G, int(d["dens_iter"]), 0.9 * W, float(W), 2.0 * synth.scene_extent(3.0))
        assert stats == tuple(int(x) for x in d["dens_stats"])
        assert G.count == len(d["dens_out"])
        dout = G.download()
        for f in ("position", "scale", "rotation", "opacity", "sh"):
            assert np.array_equal(dout[f], d["dens_out"][f]), f
    finally:
        for o in (Dn, A, D, gti, img, R, G):
            o.close()
