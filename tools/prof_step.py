"""Small driver for ncu: a few fwd+bwd(+Adam) steps of config C2 through the C ABI (no torch, no oracle).
usage: python tools/prof_step.py [workload] [warm] [steps] [noadam] [knob=value ...]
(knobs go to gs_tune; capture the ordering kernels with side_bins=0: ncu restores device memory around every replayed pass of a
kernel, which rolls back what a kernel running on the side stream wrote in the meantime)"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
warm = int(sys.argv[2]) if len(sys.argv) > 2 else 3
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
adam = "noadam" not in sys.argv[4:]
for a in sys.argv[4:]:
    if "=" in a:
        gs.tune(a.split("=")[0], int(a.split("=")[1]))
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, synth.CONFIG_SEED[wl] ^ 0xA5A5, synth.CONFIG_S0[wl]))
R = gs.Rasterizer(ctx, n)
img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
D = gs.Grads(ctx, n); A = gs.Adam(ctx, n)
cams = [synth.make_camera(v, 64, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(warm + steps + 1)]
R.forward(Gt, cams[0], gti)
for i in range(warm + steps):
    R.forward(G, cams[i + 1], img)
    R.backward(G, D, cams[i + 1], img, gti)
    if adam:
        A.step(G, D, [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3])
ctx.sync()
print("ok pairs", R.pair_count, "passes", R.sort_passes)
