import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth
wl="C2"
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
g0 = synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl])
G = gs.Gaussians(ctx, n); G.upload(g0)
Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, synth.CONFIG_SEED[wl] ^ 0xA5A5, synth.CONFIG_S0[wl]))
R = gs.Rasterizer(ctx, n)
img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
D = gs.Grads(ctx, n); A = gs.Adam(ctx, n)
cams = [synth.make_camera(v, 64, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(12)]
for v in range(8):
    R.forward(G, cams[v], img); print("view", v, "pairs", R.pair_count)
R.forward(Gt, cams[0], gti)
for i in range(8):
    R.forward(G, cams[0], img)
    R.backward(G, D, cams[0], img, gti)
    gr = D.download(n)
    A.step(G, D, [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3])
    g1 = G.download()
    print("iter", i, "pairs", R.pair_count, "dscale max", np.abs(g1["scale"]-g0["scale"]).max(), "mean", (g1["scale"]-g0["scale"]).mean(),
          "dop mean", (g1["opacity"]-g0["opacity"]).mean(), "nan scale", np.isnan(g1["scale"]).sum(), "grad scale absmax", np.abs(gr["scale"]).max(), "nonzero frac", (gr["scale"]!=0).mean())
