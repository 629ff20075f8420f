"""The C++ host layer (gaussiansplatting_b200/host/: the reference's class interfaces + the train loop of
mtl_engine.mm:1047-1221) on the GPU: the GPURadixSort64 class against a stable CPU sort, and a short training run
whose result must equal the same schedule driven step by step through the C ABI from Python."""
import numpy as np
import pytest
from gaussiansplatting_b200 import synth

pytestmark = pytest.mark.gpu


def test_cxx_sort64_class(gs):
    from gaussiansplatting_b200 import host
    n = 50000
    keys = synth.splitmix64(3, n, 1) >> np.uint64(20)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = host.sort64(keys, vals)
    o = np.argsort(keys, kind="stable")
    assert np.array_equal(ko, keys[o]) and np.array_equal(vo, vals[o])


def test_cxx_engine_train_matches_python_driver(gs, ctx):
    from gaussiansplatting_b200 import host
    n, W, H, V, orbit = 6000, 256, 160, 4, 3.0
    g = synth.make_scene(n, 71, 0.05)
    g["opacity"][::9] = -6.0                     # a few nearly transparent splats: pruned at the first densify
    gt = synth.make_scene(n, 71 ^ 0xA5A5, 0.05)
    poses = [synth.orbit_pose(v, V, orbit) for v in range(V)]
    quats = np.array([p[0] for p in poses], np.float32)
    trans = np.array([p[1] for p in poses], np.float32)
    f = 0.9 * W
    cfg = host.TrainConfig(width=W, height=H, num_views=V, epochs=6, fx=f, fy=f, cx=W / 2, cy=H / 2,
                           scene_extent=synth.scene_extent(orbit), densify_from=8, densify_until=10 ** 6, densify_interval=5,
                           opacity_reset_interval=12, seed=7, quantize_rgba8=0)
    res, out = host.train(g, gt, quats, trans, cfg, want_output=True)
    iters = cfg.epochs * V
    assert res["iterations"] == iters and res["adam_timestep"] == iters
    assert res["densify_events"] == len([i for i in range(1, iters + 1) if i > 8 and i % 5 == 0])
    assert res["opacity_resets"] == iters // 12
    assert res["final_count"] == len(out) and res["pruned"] + res["cloned"] + res["split"] > 0 and res["final_count"] != n
    # the same schedule through the C ABI from Python
    G = gs.Gaussians(ctx, n); G.upload(g)
    Gt = gs.Gaussians(ctx, n); Gt.upload(gt)
    R = gs.Rasterizer(ctx, n); D = gs.Grads(ctx, n); A = gs.Adam(ctx, n); Dn = gs.Density(ctx)
    Dn.set_scene_extent(synth.scene_extent(orbit)); Dn.set_seed(7)
    img = gs.Image(ctx, W, H)
    cams = [synth.build_uniforms(quats[v], trans[v], f, f, W / 2, H / 2, W, H, n) for v in range(V)]
    gts = [gs.Image(ctx, W, H) for _ in range(V)]
    for v in range(V):
        R.forward(Gt, cams[v], gts[v])
    it = 0
    try:
        for ep in range(cfg.epochs):
            for v in range(V):
                lr = 1.6e-4 * (1.6e-6 / 1.6e-4) ** (it / iters)
                R.forward(G, cams[v], img); R.backward(G, D, cams[v], img, gts[v])
                Dn.accumulate(D, G.count)
                A.step(G, D, [np.float32(lr), 5e-3, 1e-3, 2.5e-2, 2.5e-3])
                it += 1
                if it > 8 and it % 5 == 0:
                    old = G.count
                    Dn.apply(G, it, f, float(W), 2.0 * synth.scene_extent(orbit))
                    A.resize_if_needed(G.count)
                    if G.count > old:
                        A.reset_state_for_new(old)
                if it % 12 == 0:
                    G.clamp_opacity(-4.6); A.reset_opacity_momentum(); A.reset_scale_momentum(); Dn.reset_accumulator(G.count)
        ref = G.download()
    finally:
        for o in gts + [img, Dn, A, D, R, Gt, G]:
            o.close()
    # same kernels, same order: identical counts; values agree to the noise of the backward's float atomics
    assert len(ref) == len(out)
    for fld in ("position", "scale", "rotation", "opacity", "sh"):
        a, b = out[fld].astype(np.float64), ref[fld].astype(np.float64)
        assert np.quantile(np.abs(a - b), 0.999) <= 1e-3, fld


def test_cxx_engine_logging_loss_and_ppm_export(gs, tmp_path):
    """Engine::computeLoss in every trainStep (the reference's behaviour, mtl_engine.mm:991), exportTrainingViews /
    saveTextureToPPM from a headless device image (mtl_engine.mm:19-63, 1224-1306)"""
    from gaussiansplatting_b200 import host
    n, W, H, V, orbit = 4000, 128, 96, 3, 3.0
    g = synth.make_scene(n, 81, 0.05)
    gt = synth.make_scene(n, 81 ^ 0xA5A5, 0.05)
    poses = [synth.orbit_pose(v, V, orbit) for v in range(V)]
    quats = np.array([p[0] for p in poses], np.float32)
    trans = np.array([p[1] for p in poses], np.float32)
    f = 0.9 * W
    cfg = host.TrainConfig(width=W, height=H, num_views=V, epochs=20, fx=f, fy=f, cx=W / 2, cy=H / 2,
                           scene_extent=synth.scene_extent(orbit), densify_from=10 ** 6, densify_until=10 ** 6 + 1, densify_interval=5,
                           opacity_reset_interval=10 ** 6, seed=7, quantize_rgba8=1, log_loss_every_step=1,
                           export_dir=str(tmp_path).encode())
    res, _ = host.train(g, gt, quats, trans, cfg)
    assert 0.0 < res["last_loss"] < res["first_loss"] < 1.0          # 60 Adam steps reduce the logging loss
    assert res["views_exported"] == V
    for v in range(V):
        raw = (tmp_path / f"view_{v:03d}.ppm").read_bytes()
        head = f"P6\n{W} {H}\n255\n".encode()
        assert raw.startswith(head) and len(raw) == len(head) + W * H * 3
    px = np.frombuffer(raw[len(head):], np.uint8)
    assert px.min() < 250 and px.max() == 255                       # splats over the white background


def test_cxx_engine_render_matches_the_viewer_oracle(gs, orc):
    """Engine::render (MTLEngine::render, mtl_engine.mm:401-609, headless): GPURadixSort32 draw order + the instanced-quad
    draw through the C++ classes, against the CPU restatement of the viewer shaders"""
    from gaussiansplatting_b200 import host
    n, W, H = 3000, 160, 112
    g = synth.make_scene(n, 41, 0.04)
    q, t = synth.orbit_pose(2, 8, 3.0)
    f = 0.9 * W
    img, pairs = host.render_view(g, q, t, f, f, W / 2.0, H / 2.0, W, H)
    u = synth.make_camera(2, 8, W, H, n)
    vu = orc.view_uniforms(u)
    order = orc.viewer_order(g, u["cameraPos"][0][:3])
    want, shaded, blended = orc.viewer_render(g, order, vu, threads=0)
    sp = orc.viewer_splats(g, vu)
    v = sp[sp["valid"] == 1]
    assert pairs == int(((v["tileMaxX"] - v["tileMinX"] + 1) * (v["tileMaxY"] - v["tileMinY"] + 1)).sum())
    assert blended > 0 and np.abs(img - want).max() <= 1e-4
    img8, _ = host.render_view(g, q, t, f, f, W / 2.0, H / 2.0, W, H, rgba8_target=True)
    want8 = orc.viewer_render(g, order, vu, quantize=True, threads=0)[0]
    assert np.abs(img8 - want8).max() <= 1e-4
