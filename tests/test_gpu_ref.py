"""The CUDA path against the REFERENCE'S OWN OBJECT CODE (oracle/_ref/libref.so, compiled from /root/reference sources:
oracle/Makefile) for the halves of the hot path the reference runs on the CPU:
  gs_sort64_sort          <-> parallelRadixSort                       tiled_rasterizer.mm:27-102 (what ships: USE_GPU_SORT = false)
  gs_dens_accumulate      <-> DensityController::accumulateGradients   density_control.mm:121-185
  gs_dens_apply           <-> DensityController::apply                 density_control.mm:188-501
  the frame's sorted list <-> parallelRadixSort of the frame's own pre-sort pairs (generateTilePairs output)
"""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth
from oracle import ref as R

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref.so absent (it is built where /root/reference exists)")]


@pytest.mark.parametrize("n", [1000, 4097, 1 << 16, 1 << 20, (1 << 22) + 12345])
def test_gpu_sort64_is_the_references_cpu_sort(gs, ctx, n):
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
    keys[: n // 2] = (rng.integers(0, 8160, n // 2, dtype=np.uint64) << np.uint64(32)) | (rng.integers(0, 256, n // 2, dtype=np.uint64) | np.uint64(1 << 31))
    vals = rng.permutation(n).astype(np.uint32)
    rk, rv = R.sort_pairs(keys, vals)
    S = gs.Sort64(ctx, n)
    try:
        gk, gv = S.sort_host(keys, vals)
    finally:
        S.close()
    assert np.array_equal(gk, rk) and np.array_equal(gv, rv)


def test_frame_lists_are_the_references_sort_of_the_frames_pairs(gs, ctx):
    """the per-tile lists of a C1 frame (binned pipeline: no pair sort runs at all) are, element for element, what the
    reference's CPU sort makes of the frame's pre-sort (key, value) list"""
    n, W, H = 10000, 256, 256
    G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED["C1"], synth.CONFIG_S0["C1"]))
    Rz = gs.Rasterizer(ctx, n); img = gs.Image(ctx, W, H)
    try:
        Rz.forward(G, synth.make_camera(2, 8, W, H, n), img)
        keys, vals = Rz.get(gs.RAST_KEYS), Rz.get(gs.RAST_VALUES)
        assert len(keys) > 20000
        rk, rv = R.sort_pairs(keys, vals)
        assert np.array_equal(Rz.get(gs.RAST_SORTED_KEYS), rk) and np.array_equal(Rz.get(gs.RAST_SORTED_VALUES), rv)
    finally:
        img.close(); Rz.close(); G.close()


def test_gpu_accumulate_gradients_is_the_references(gs, ctx):
    from oracle.orc import GRADIENTS
    n = 6007
    acc_r, cnt_r = np.zeros(n, np.float32), np.zeros(n, np.uint32)
    D = gs.Grads(ctx, n); Dn = gs.Density(ctx)
    try:
        Dn.reset_accumulator(n)
        for step in range(3):
            rng = np.random.default_rng(90 + step)
            g = np.zeros(n, GRADIENTS)
            g["viewspace"] = (rng.standard_normal((n, 2)) * 1e-3).astype(np.float32)
            g["viewspace"][::11] = 0; g["viewspace"][3::17, 0] = np.nan; g["viewspace"][5::19, 1] = np.inf; g["viewspace"][7::23] = 40.0
            R.accumulate_gradients(g, acc_r, cnt_r)
            D.upload(g)
            Dn.accumulate(D, n)
        acc, cnt = Dn.download(n)
        assert np.array_equal(cnt, cnt_r) and np.array_equal(acc.view(np.uint32), acc_r.view(np.uint32))
    finally:
        Dn.close(); D.close()


@pytest.mark.parametrize("iteration", [400, 700, 3500])
def test_gpu_densify_apply_is_the_references(gs, ctx, iteration):
    """markers / statistics / count / layout exact; every kept and cloned row bit-equal; split children: scale, rotation,
    opacity and SH bit-equal, positions symmetric about the parent with a unit direction in the parent's scaled frame (the
    reference draws the direction from rand(), this library from a counter-based generator: SURVEY note L)"""
    n = 5000
    g = synth.make_scene(n, 17, 0.05)
    rng = np.random.default_rng(17)
    g["opacity"][::7] = -7.0
    g["scale"][5::11] = np.log(0.5)
    counts = rng.integers(0, 4, n).astype(np.uint32)
    accum = (rng.random(n) * 1e-3 * np.maximum(counts, 1)).astype(np.float32)
    accum[::2] *= 5.0
    extent, focal, W = synth.scene_extent(3.0), 460.0, 512.0
    rout, rmark, rstats = R.densify_apply(g, accum, counts, iteration, extent, focal, W, 2 * extent, 99)
    G = gs.Gaussians(ctx, 2 * n); G.upload(g)
    Dn = gs.Density(ctx)
    try:
        Dn.set_scene_extent(extent); Dn.set_seed(0x5EED)
        Dn.upload(accum, counts)
        stats = Dn.apply(G, iteration, focal, W, 2 * extent)
        _, _, mark = Dn.download(n, markers=True)
        out = G.download()
    finally:
        Dn.close(); G.close()
    assert stats == tuple(int(x) for x in rstats) and np.array_equal(mark, rmark) and len(out) == len(rout)
    src = np.repeat(np.arange(n), np.where(rmark == 1, 0, np.where(rmark == 0, 1, 2)))
    child = rmark[src] == 3
    for f in ("scale", "rotation", "opacity", "sh"):
        assert np.array_equal(out[f].view(np.uint32), rout[f].view(np.uint32)), f
    assert np.array_equal(out["position"][~child].view(np.uint32), rout["position"][~child].view(np.uint32))
    if child.any():
        assert iteration > 500
        c = out["position"][child].astype(np.float64).reshape(-1, 2, 3)
        par = g["position"][src[child]].astype(np.float64).reshape(-1, 2, 3)[:, 0]
        assert np.abs((c[:, 0] + c[:, 1]) / 2 - par).max() < 1e-6
        q = g["rotation"][src[child]].reshape(-1, 2, 4)[:, 0].astype(np.float64)
        w, x, y, z = q.T
        Rm = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], -1),
                       np.stack([2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)], -1),
                       np.stack([2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)], -2)   # [row][col]
        local = np.einsum("nij,nj->ni", np.linalg.inv(Rm), c[:, 0] - par)
        sc = np.exp(np.clip(g["scale"][src[child]].reshape(-1, 2, 3)[:, 0].astype(np.float64), -4, 4))
        assert np.abs(np.linalg.norm(local / sc, axis=1) - 1).max() < 1e-4
