"""The reference's training loop (mtl_engine.mm:856-1025 trainStep, :1047-1221 train) driven over the CPU oracle alone:
forward -> backward -> accumulateGradients -> Adam -> (densify -> resizeIfNeeded -> resetStateForNewGaussians, SURVEY note D)
-> opacity reset, with the reference's own schedule constants.  Test helper (not collected by pytest)."""
import ctypes
import os
import numpy as np
from gaussiansplatting_b200 import synth

LRS = (5e-3, 1e-3, 2.5e-2, 2.5e-3)                    # scale, rotation, opacity, sh: mtl_engine.mm:1063-1068
WIDTHS = (3, 3, 4, 1, 12)                             # position, scale, rotation, opacity, sh moment widths (optimizer.mm:34-60)


def run(orc, g0, gt, quats, trans, W, H, f, extent, epochs, start, total, densify_interval, seed, capacity):
    """returns (final Gaussians, events dict, AdamState)"""
    th = len(os.sched_getaffinity(0))
    n0, V = len(g0), len(quats)
    cams = [synth.build_uniforms(quats[v], trans[v], f, f, W / 2, H / 2, W, H, n0) for v in range(V)]
    gts = [orc.render(gt, cams[v], th)["image"] for v in range(V)]
    st = orc.AdamState(capacity)                       # the reference sizes Adam for 2,000,000 (mtl_engine.mm:318)
    accum, counts = np.zeros(capacity, np.float32), np.zeros(capacity, np.uint32)
    g = np.ascontiguousarray(g0.copy())
    it = start
    ev = dict(densify=0, resets=0, pruned=0, cloned=0, split=0)
    decay = orc.lib().orc_lr_decay
    for ep in range(epochs):
        for v in range(V):
            n = len(g)
            assert n <= capacity
            u = cams[v].copy(); u["numGaussians"] = n
            lr_pos = float(decay(ctypes.c_float(1.6e-4), ctypes.c_float(1.6e-6), ctypes.c_uint64(it), ctypes.c_uint64(total)))   # :1090
            r = orc.render(g, u, th)                                                                   # :973
            gr = orc.backward(g, r["proj"], r["svals"], r["ranges"], u, r["last"], r["image"], gts[v], threads=th)   # :994
            a, c = accum[:n].copy(), counts[:n].copy()
            orc.accumulate_gradients(gr, a, c)                                                         # :998
            accum[:n], counts[:n] = a, c
            # Adam over the live count (note A) on exactly-n-long moment arrays
            m = [np.ascontiguousarray(st.m[k][:n * w]) for k, w in enumerate(WIDTHS)]
            vv = [np.ascontiguousarray(st.v[k][:n * w]) for k, w in enumerate(WIDTHS)]
            live = orc.AdamState(0); live.m, live.v, live.t = m, vv, st.t
            orc.adam_step(g, gr, live, [lr_pos, *LRS], threads=th)                                      # :1001
            st.t = live.t
            for k, w in enumerate(WIDTHS):
                st.m[k][:n * w] = m[k]; st.v[k][:n * w] = vv[k]
            it += 1
            if 500 < it < 15000 and it % densify_interval == 0:                                        # :1112-1168
                g, _, ds = orc.densify_apply(g, accum[:n].copy(), counts[:n].copy(), it, extent, f, float(W), 2.0 * extent, seed)
                g = np.ascontiguousarray(g)
                accum[:len(g)] = 0; counts[:len(g)] = 0          # resetAccumulator(gaussianCount), density_control.mm:493
                if len(g) > n:                                    # resetStateForNewGaussians(oldCount): zero [old, capacity)
                    for k, w in enumerate(WIDTHS):
                        st.m[k][n * w:] = 0; st.v[k][n * w:] = 0
                ev["densify"] += 1; ev["pruned"] += int(ds[0]); ev["cloned"] += int(ds[1]); ev["split"] += int(ds[2])
            if it % 3000 == 0 and 0 < it < 15000:                                                      # :1173-1192
                g["opacity"] = np.minimum(g["opacity"], np.float32(-4.6))
                st.m[3][:] = 0; st.v[3][:] = 0; st.m[1][:] = 0; st.v[1][:] = 0
                accum[:len(g)] = 0; counts[:len(g)] = 0
                ev["resets"] += 1
    return g, ev, st
