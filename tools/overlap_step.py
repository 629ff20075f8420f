"""The overlap mode inside ONE training step of several views (BASELINE config 5 on one GPU: 8 views per step, the parameters are
fixed until the step's Adam): two rasterizers alternate over the views, gs_rast_set_overlap re-armed after every Adam step.
usage: python tools/overlap_step.py [views_per_step] [steps]"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussiansplatting_b200 import capi as gs, synth

vps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
wl = "C5"
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
LRS = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]
ctx = gs.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
g0 = synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl])
G = gs.Gaussians(ctx, n)
Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, synth.CONFIG_SEED[wl] ^ 0xA5A5, synth.CONFIG_S0[wl]))
Rs = [gs.Rasterizer(ctx, n), gs.Rasterizer(ctx, n)]
img = gs.Image(ctx, W, H)
cams = [synth.make_camera(v, 16, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(16)]
gts = []
for v in range(16):
    t = gs.Image(ctx, W, H)
    for R in Rs:
        R.forward(Gt, cams[v], t)              # also sizes both rasterizers' pair buffers
    gts.append(t)
ctx.sync(); Gt.close()
D = gs.Grads(ctx, n); A = gs.Adam(ctx, n)
for ov in (0, 1, 0, 1):
    G.upload(g0); A.reset()
    for R in Rs:
        R.set_overlap(bool(ov))

    def step(s):
        for k in range(vps):
            v = (s * vps + k) % 16
            R = Rs[k % 2]
            R.forward(G, cams[v], img)
            R.backward(G, D, cams[v], img, gts[v], gs.GS_BWD_ACCUMULATE if k else 0)
        A.step(G, D, LRS)
        if ov:
            for R in Rs:
                R.set_overlap(True)            # re-arm: the next forwards must wait for this Adam step
    for s in range(3):
        step(s)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for s in range(steps):
        step(s)
    e1.record(stream); e1.synchronize()
    p = G.download()
    print(f"C5 step of {vps} views, overlap={ov}: {e0.elapsed_time(e1) / steps:.3f} ms/step, {1e3 * steps / e0.elapsed_time(e1):.1f} steps/s, "
          f"param checksum {float(np.abs(p['position']).sum()):.6e}", flush=True)
