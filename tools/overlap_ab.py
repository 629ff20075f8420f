"""Two rasterizers alternating over independent frames, with and without gs_rast_set_overlap (the ordering stage of frame i+1 on a
high-priority side stream under the blend of frame i).  usage: python tools/overlap_ab.py [workload] [frames] [render|train]"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussiansplatting_b200 import capi as gs, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 32
mode = sys.argv[3] if len(sys.argv) > 3 else "train"
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
Rs = [gs.Rasterizer(ctx, n), gs.Rasterizer(ctx, n)]
imgs = [gs.Image(ctx, W, H), gs.Image(ctx, W, H)]
gti = gs.Image(ctx, W, H)
D = gs.Grads(ctx, n)
cams = [synth.make_camera(v, 64, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(64)]
if mode == "train":
    Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, synth.CONFIG_SEED[wl] ^ 0xA5A5, synth.CONFIG_S0[wl]))
    Rs[0].forward(Gt, cams[0], gti); ctx.sync(); Gt.close()


def frame(i):
    R, img = Rs[i % 2], imgs[i % 2]
    R.forward(G, cams[i % 64], img)
    if mode == "train":
        R.backward(G, D, cams[i % 64], img, gti)


for R in Rs:                                   # size the pair buffers of both
    for v in range(0, 64, 1):
        R.forward(G, cams[v], imgs[0])
ctx.sync()
ref = {}
for rep in range(2):
    for ov in (0, 1):
        for R in Rs:
            R.set_overlap(bool(ov))
        for i in range(4):
            frame(i)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(frames):
            frame(i)
        e1.record(stream); e1.synchronize()
        ms = e0.elapsed_time(e1) / frames
        ctx.sync()
        last = (frames - 1) % 2
        chk = float(np.abs(imgs[last].download()).sum())
        if mode == "train":
            g = D.download(n)
            chk += float(np.abs(g["position"]).sum() + np.abs(g["scale"]).sum())
        ref.setdefault("chk", chk)
        print(f"{wl} {mode} overlap={ov}: {ms:.4f} ms/frame  checksum rel diff {abs(chk - ref['chk']) / ref['chk']:.2e}", flush=True)
