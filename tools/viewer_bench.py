"""Viewer frame (gs_sort32_sort_depth + gs_viewer_render) on a workload's scene: device time per frame.
usage: python tools/viewer_bench.py [workload] [reps]"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth
from gaussiansplatting_b200.layouts import view_uniforms
import torch

wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
S, V, img = gs.Sort32(ctx, n), gs.Viewer(ctx, n), gs.Image(ctx, W, H)
pos = ctx.alloc(n * 16); G.positions(pos)
cams = [synth.make_camera(v, 64, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(8)]


def frame(i):
    u = cams[i % 8]
    S.sort_depth(pos, u["cameraPos"][0][:3], n)
    V.render(G, S.sorted_indices, view_uniforms(u), img)


for rep in range(2):
    for order in (0, 1):                       # draw CTAs in raster order / longest tile list first (default)
        gs.tune("tile_order", order)
        for i in range(3):
            frame(i)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(reps):
            frame(i)
        e1.record(stream); e1.synchronize()
        print(f"{wl} tile_order={order}: {e0.elapsed_time(e1) / reps:.3f} ms / viewer frame, {V.pair_count} pairs")
