"""device time of gs_adam_step alone at 1 M Gaussians with the gradients of a real C5-scene frame (CUDA events, 50 calls);
usage: python tools/adam_bench.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussiansplatting_b200 import capi as gs, synth

wl = "C5"
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
Gt = gs.Gaussians(ctx, n); Gt.upload(synth.make_scene(n, synth.CONFIG_SEED[wl] ^ 0xA5A5, synth.CONFIG_S0[wl]))
R = gs.Rasterizer(ctx, n); D = gs.Grads(ctx, n); A = gs.Adam(ctx, n)
img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
u = synth.make_camera(1, 16, W, H, n, synth.CONFIG_ORBIT[wl])
R.forward(Gt, u, gti); R.forward(G, u, img); R.backward(G, D, u, img, gti)
lrs = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]
for _ in range(5):
    A.step(G, D, lrs)
ctx.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
# a buffer larger than L2 is written between calls so that every step streams from HBM
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
tot = 0.0
for _ in range(30):
    with torch.cuda.stream(stream):
        flush.zero_()
    e0.record(stream); A.step(G, D, lrs); e1.record(stream); e1.synchronize()
    tot += e0.elapsed_time(e1)
print(f"adam step, {n} Gaussians, L2 flushed between calls: {tot / 30 * 1e3:.1f} us")
e0.record(stream)
for _ in range(50):
    A.step(G, D, lrs)
e1.record(stream); e1.synchronize()
print(f"adam step, back to back: {e0.elapsed_time(e1) / 50 * 1e3:.1f} us")
