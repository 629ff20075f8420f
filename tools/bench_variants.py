"""Per-stage device times of fwd+bwd on a workload for each blend-kernel variant (pixels per thread).
usage: python tools/bench_variants.py [workload] [reps]"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, synth.CONFIG_SEED[wl] ^ 0xA5A5, synth.CONFIG_S0[wl]))
R = gs.Rasterizer(ctx, n)
img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
D = gs.Grads(ctx, n)
cams = [synth.make_camera(v, 64, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(8)]
R.forward(Gt, cams[0], gti)
R.set_profiling(True)
ref = None
for mode, fp, bp in ((0, 2, 2), (2, 2, 2), (1, 2, 2)):
    gs.tune("sort_mode", mode); gs.tune("fwd_ppt", fp); gs.tune("bwd_ppt", bp)
    acc = None
    for i in range(3 + reps):
        R.forward(G, cams[i % 8], img)
        R.backward(G, D, cams[i % 8], img, gti)
        t = R.timings()
        if i >= 3:
            acc = t if acc is None else {k: acc[k] + t[k] for k in t}
    st = {k: v / reps for k, v in acc.items()}
    R.forward(G, cams[0], img); R.backward(G, D, cams[0], img, gti)
    g = D.download(n)
    chk = float(np.abs(g["position"]).sum() + np.abs(g["scale"]).sum())
    if ref is None:
        ref = chk
    print(f"mode {mode} fwd_ppt {fp} bwd_ppt {bp}: total {sum(st.values()):.3f} ms | " + " ".join(f"{k} {v*1e3:.0f}" for k, v in st.items())
          + f" | grad checksum rel diff {abs(chk-ref)/ref:.2e}")
print("pairs", R.pair_count, "E,C", R.count_evals())
