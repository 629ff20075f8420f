"""Small driver for ncu: a few render-only frames through the C ABI.  usage: python tools/prof_render.py [workload] [frames] [knob=value ...]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "C4"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for a in sys.argv[3:]:
    if "=" in a:
        gs.tune(a.split("=")[0], int(a.split("=")[1]))
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
R = gs.Rasterizer(ctx, n)
img = gs.Image(ctx, W, H)
for i in range(frames):
    R.forward(G, synth.make_camera(i, 64, W, H, n, synth.CONFIG_ORBIT[wl]), img)
ctx.sync()
print("ok pairs", R.pair_count, "coarse", R.coarse_pair_count, "passes", R.sort_passes)
