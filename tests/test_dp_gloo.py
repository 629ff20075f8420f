"""Host logic of the N > 1 path on CPU: two gloo ranks deal the views of a step, accumulate per-view packed gradients
(from the oracle, standing in for the CUDA backward), all-reduce once, and must end with the single-process sum and
with bit-identical parameters after the Adam step on every rank."""
import os
import socket
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_view_dealing():
    from gaussiansplatting_b200 import dp
    for world in (1, 2, 4, 8):
        for step in (0, 1, 7):
            dealt = [dp.views_for_rank(step, 8, world, r, 64) for r in range(world)]
            flat = sorted(v for d in dealt for v in d)
            assert flat == sorted((step * 8 + s) % 64 for s in range(8))          # a partition of the step's views
            assert all(len(d) == 8 // world for d in dealt)                       # balanced
        got = sorted(v for r in range(world) for v in dp.render_views_for_rank(13, world, r))
        assert got == list(range(13))


def test_pack_round_trip(orc):
    from gaussiansplatting_b200 import dp
    n = 100
    g = np.zeros(n, orc.GRADIENTS)
    r = np.random.RandomState(0).randn(n, 16).astype(np.float32)
    g["position"], g["opacity"], g["scale"], g["rotation"] = r[:, :3], r[:, 3], r[:, 4:7], r[:, 7:11]
    g["sh"][:, [0, 4, 8]] = r[:, 11:14]
    g["viewspace"] = r[:, 14:16]
    back = dp.unpack_gradients(dp.pack_gradients(g, False), orc.GRADIENTS)
    for f in ("position", "opacity", "scale", "rotation", "sh", "viewspace"):
        assert np.array_equal(back[f], g[f]), f
    p = dp.pack_gradients(g, True)
    assert np.all(p[15] == 1) and np.allclose(p[14], np.minimum(np.hypot(r[:, 14], r[:, 15]), 1), rtol=1e-6)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from gaussiansplatting_b200 import synth, dp
    from oracle import orc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, W, H, V = 400, 64, 48, 4
        g = synth.make_scene(n, 21, 0.05)
        gt_g = synth.make_scene(n, 21 ^ 0xA5A5, 0.05)
        vp = dp.ViewParallel(dist)
        packed = np.zeros((dp.PLANES, n), np.float32)
        for v in vp.views(0, V, 8):
            u = synth.make_camera(v, 8, W, H, n)
            r = orc.render(g, u)
            gt = orc.render(gt_g, u)["image"]
            gr = orc.backward(g, r["proj"], r["svals"], r["ranges"], u, r["last"], r["image"], gt)
            packed += dp.pack_gradients(gr, True)
        t = torch.from_numpy(packed)
        vp.allreduce_(t)
        grads = dp.unpack_gradients(t.numpy(), orc.GRADIENTS)
        st = orc.AdamState(n)
        orc.adam_step(g, grads, st, [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3])
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), packed=t.numpy(), g=g)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_allreduce_matches_single_process(orc, tmp_path):
    import torch.multiprocessing as mp
    from gaussiansplatting_b200 import synth, dp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    assert np.array_equal(a["packed"], b["packed"])                 # every rank holds the same reduced buffer
    for f in ("position", "scale", "rotation", "opacity", "sh"):
        assert np.array_equal(a["g"][f], b["g"][f]), f              # identical Adam step on every rank
    # single process, all four views
    n, W, H, V = 400, 64, 48, 4
    g = synth.make_scene(n, 21, 0.05)
    gt_g = synth.make_scene(n, 21 ^ 0xA5A5, 0.05)
    ref = np.zeros((dp.PLANES, n), np.float64)
    for v in dp.views_for_rank(0, V, 1, 0, 8):
        u = synth.make_camera(v, 8, W, H, n)
        r = orc.render(g, u)
        gt = orc.render(gt_g, u)["image"]
        gr = orc.backward(g, r["proj"], r["svals"], r["ranges"], u, r["last"], r["image"], gt)
        ref += dp.pack_gradients(gr, True)
    got = a["packed"].astype(np.float64)
    scale = np.abs(ref).max(axis=1, keepdims=True) + 1e-30
    assert np.all(np.abs(got - ref) <= 1e-5 * np.maximum(np.abs(ref), np.abs(got)) + 1e-6 * scale)
    assert np.array_equal(got[15], ref[15])                         # view counts are integers: exact
