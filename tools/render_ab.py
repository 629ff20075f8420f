"""A/B of one gs_tune knob on the render-only frame (forward alone; no backward has run on the rasterizer, so no visit words are
recorded).  usage: python tools/render_ab.py KNOB [workload] [frames] [values]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussiansplatting_b200 import capi as gs, synth

knob = sys.argv[1]
wl = sys.argv[2] if len(sys.argv) > 2 else "C4"
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 20
values = [int(x) for x in sys.argv[4].split(",")] if len(sys.argv) > 4 else [0, 2]
n = synth.CONFIG_N[wl]; W, H = synth.CONFIG_RES[wl]
ctx = gs.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, synth.CONFIG_SEED[wl], synth.CONFIG_S0[wl]))
R = gs.Rasterizer(ctx, n)
img = gs.Image(ctx, W, H)
cams = [synth.make_camera(v, 64, W, H, n, synth.CONFIG_ORBIT[wl]) for v in range(64)]
for kv in os.environ.get("GS_TUNE", "").split():       # fixed knobs for the whole run, e.g. GS_TUNE="fwd_warp=2"
    gs.tune(kv.split("=")[0], int(kv.split("=")[1]))
for rep in range(3):
    for v in values:
        gs.tune(knob, v)
        for i in range(3):
            R.forward(G, cams[i], img)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(frames):
            R.forward(G, cams[i % 64], img)
        e1.record(stream); e1.synchronize()
        ms = e0.elapsed_time(e1) / frames
        R.set_profiling(True)
        acc = {}
        for i in range(frames):
            R.forward(G, cams[i % 64], img); ctx.sync()
            for k, t in R.timings(False).items():
                acc[k] = acc.get(k, 0.0) + t / frames
        R.set_profiling(False)
        print(f"{knob}={v}: {ms:.4f} ms/frame | " + " ".join(f"{k} {t * 1e3:.0f}" for k, t in acc.items() if t > 0), flush=True)
